// lr.cu — sparse logistic regression iterations with the parameter-server traffic fused in.
//
// Reference hot loop (apps/logistic_regression/lr_example.cpp:247-315):
//     table->Get(keys,&params); for sample in batch: pred = sigmoid(sum w_j x_j);
//     delta_j += alpha*x_j*(y-pred); table->Add(keys,deltas); table->Clock();
// with Get/Add each a ZMQ round trip through a server thread.  Here the pull, the forward
// SpMV, the sigmoid/loss, the gradient, the push and (dense path) the clock edge are
// phases of one kernel that addresses the owning shards directly over NVLink.
//
// Two regimes:
//   sparse  (batch nnz << table size, e.g. the 100M-key model): per-nnz peer ld / red.add.
//   dense   (batch touches most of the table, e.g. 64K rows x 40 nnz over 1M keys):
//           slab-pull every shard into an L2-resident replica with 16-byte peer loads,
//           accumulate the batch gradient locally (the device form of the reference's
//           per-batch unique-key pre-aggregation, lib/batch_data_sampler.cpp:50-71), then
//           slab-push with vector reds; whole iterations loop inside a persistent
//           cooperative kernel so no host round trip gates an iteration.
#include <cooperative_groups.h>

#include <cstdio>
#include <mutex>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace mps {

__device__ __forceinline__ float sigmoidf_fast(float z) { return 1.f / (1.f + __expf(-z)); }
// logloss of label y in {0,1} at logit z, numerically stable: softplus(z) - y*z
__device__ __forceinline__ float logloss(float z, float y) {
  return fmaxf(z, 0.f) + log1pf(__expf(-fabsf(z))) - y * z;
}

// block reduce of (loss, correct) pairs then one atomic pair per CTA
__device__ __forceinline__ void block_accumulate(float loss, float correct, float* out) {
  __shared__ float sl[32], sc[32];
  loss = warp_sum(loss);
  correct = warp_sum(correct);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) { sl[wid] = loss; sc[wid] = correct; }
  __syncthreads();
  if (wid == 0) {
    const int nw = (blockDim.x + 31) >> 5;
    float a = lane < nw ? sl[lane] : 0.f, b = lane < nw ? sc[lane] : 0.f;
    a = warp_sum(a);
    b = warp_sum(b);
    if (lane == 0 && (a != 0.f || b != 0.f)) {
      atomicAdd(out, a);
      atomicAdd(out + 1, b);
    }
  }
  __syncthreads();
}

// ------------------------------------------------------------------ sparse regime
__global__ void __launch_bounds__(256)
lr_step_sparse_kernel(const __grid_constant__ MpsTableView v, const __grid_constant__ MpsClockView c,
                      int gated, int progress, const __grid_constant__ MpsCsr csr, int64_t row_begin,
                      int64_t batch, float alpha, float* __restrict__ loss_out) {
  if (gated) cta_gate(c, progress);
  const int lane = threadIdx.x & 31;
  const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  float loss = 0.f, correct = 0.f;
  for (int64_t r = warp; r < batch; r += nwarps) {
    int64_t ridx = row_begin + r;
    if (ridx >= csr.n_rows) ridx %= csr.n_rows;
    const int64_t beg = csr.row_ptr[ridx], end = csr.row_ptr[ridx + 1];
    // pull + forward: two nnz per lane in flight (rows of <= 64 nnz stay in registers)
    uint32_t c0 = 0, c1 = 0;
    float x0 = 0.f, x1 = 0.f;
    float* d0 = nullptr;
    float* d1 = nullptr;
    float acc = 0.f;
    for (int64_t base = beg; base < end; base += 64) {
      const int64_t j0 = base + lane, j1 = j0 + 32;
      uint32_t kc0 = 0, kc1 = 0;
      float kx0 = 0.f, kx1 = 0.f;
      if (j0 < end) { kc0 = ld_stream(csr.col + j0); kx0 = ld_stream(csr.val + j0); }
      if (j1 < end) { kc1 = ld_stream(csr.col + j1); kx1 = ld_stream(csr.val + j1); }
      float* p0 = nullptr;
      float* p1 = nullptr;
      float w0 = 0.f, w1 = 0.f;
      if (j0 < end) {
        const int s = owner_of(v, kc0);
        if (s >= 0) {
          w0 = ld_sys(static_cast<const float*>(v.weights[s]) + (kc0 - v.begin[s]));
          p0 = static_cast<float*>(v.staging[s] ? v.staging[s] : v.weights[s]) + (kc0 - v.begin[s]);
        }
      }
      if (j1 < end) {
        const int s = owner_of(v, kc1);
        if (s >= 0) {
          w1 = ld_sys(static_cast<const float*>(v.weights[s]) + (kc1 - v.begin[s]));
          p1 = static_cast<float*>(v.staging[s] ? v.staging[s] : v.weights[s]) + (kc1 - v.begin[s]);
        }
      }
      acc = fmaf(w0, kx0, acc);
      acc = fmaf(w1, kx1, acc);
      if (base == beg) { c0 = kc0; c1 = kc1; x0 = kx0; x1 = kx1; d0 = p0; d1 = p1; }
    }
    (void)c0; (void)c1;
    const float z = warp_sum(acc);
    const float y = csr.label[ridx];
    const float pred = sigmoidf_fast(z);
    const float g = alpha * (y - pred);
    if (lane == 0) {
      loss += logloss(z, y);
      correct += ((pred > 0.5f) == (y > 0.5f)) ? 1.f : 0.f;
    }
    // gradient + push: red.add at the owner (server/vector_storage.hpp:28-38)
    if (d0 && x0 != 0.f) red_sys(d0, g * x0);
    if (d1 && x1 != 0.f) red_sys(d1, g * x1);
    for (int64_t base = beg + 64; base < end; base += 64) {
      const int64_t j0 = base + lane, j1 = j0 + 32;
      if (j0 < end) {
        const uint32_t kc = ld_stream(csr.col + j0);
        const float kx = ld_stream(csr.val + j0);
        const int s = owner_of(v, kc);
        if (s >= 0 && kx != 0.f)
          red_sys(static_cast<float*>(v.staging[s] ? v.staging[s] : v.weights[s]) + (kc - v.begin[s]), g * kx);
      }
      if (j1 < end) {
        const uint32_t kc = ld_stream(csr.col + j1);
        const float kx = ld_stream(csr.val + j1);
        const int s = owner_of(v, kc);
        if (s >= 0 && kx != 0.f)
          red_sys(static_cast<float*>(v.staging[s] ? v.staging[s] : v.weights[s]) + (kc - v.begin[s]), g * kx);
      }
    }
  }
  block_accumulate(loss, correct, loss_out);
}

// ------------------------------------------------------------------ dense regime
__device__ __forceinline__ int64_t flat_begin1(const MpsTableView& v, int s) {
  return static_cast<int64_t>(v.begin[s] - v.begin[0]);
}

// slab-pull all shards into replica[0 .. num_dims)
__device__ __forceinline__ void slab_pull(const MpsTableView& v, int rot_shard, float* __restrict__ replica,
                                          int64_t tid, int64_t nthreads) {
  const int S = v.num_shards;
  for (int i = 0; i < S; ++i) {
    int s = rot_shard + i;
    if (s >= S) s -= S;
    const int64_t fb = flat_begin1(v, s);
    const int64_t len = static_cast<int64_t>(v.begin[s + 1] - v.begin[s]);
    const float* src = static_cast<const float*>(v.weights[s]);
    float* dst = replica + fb;
    if ((fb & 3) == 0) {
      const int64_t len4 = len >> 2;
      for (int64_t q = tid; q < len4; q += nthreads)
        *reinterpret_cast<float4*>(dst + 4 * q) = ld_sys_v4(src + 4 * q);
      for (int64_t q = (len4 << 2) + tid; q < len; q += nthreads) dst[q] = ld_sys(src + q);
    } else {
      for (int64_t q = tid; q < len; q += nthreads) dst[q] = ld_sys(src + q);
    }
  }
}

// slab-push gacc into the owners (vector reds), clearing gacc behind itself
__device__ __forceinline__ void slab_push(const MpsTableView& v, int rot_shard, float* __restrict__ gacc,
                                          int64_t tid, int64_t nthreads) {
  const int S = v.num_shards;
  for (int i = 0; i < S; ++i) {
    int s = rot_shard + i;
    if (s >= S) s -= S;
    const int64_t fb = flat_begin1(v, s);
    const int64_t len = static_cast<int64_t>(v.begin[s + 1] - v.begin[s]);
    float* dst = static_cast<float*>(v.staging[s] ? v.staging[s] : v.weights[s]);
    float* src = gacc + fb;
    if ((fb & 3) == 0) {
      const int64_t len4 = len >> 2;
      for (int64_t q = tid; q < len4; q += nthreads) {
        const float4 x = ld_cg_v4(src + 4 * q);
        if (x.x != 0.f || x.y != 0.f || x.z != 0.f || x.w != 0.f) {
          red_sys_v4(dst + 4 * q, x);
          *reinterpret_cast<float4*>(src + 4 * q) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
      for (int64_t q = (len4 << 2) + tid; q < len; q += nthreads) {
        const float x = ld_cg(src + q);
        if (x != 0.f) { red_sys(dst + q, x); src[q] = 0.f; }
      }
    } else {
      for (int64_t q = tid; q < len; q += nthreads) {
        const float x = ld_cg(src + q);
        if (x != 0.f) { red_sys(dst + q, x); src[q] = 0.f; }
      }
    }
  }
}

// One row of the batch against a local dense copy of the weights.  kTrain: also scatter the
// gradient into gacc (device-scope red into the local L2).
template <bool kTrain>
__device__ __forceinline__ void row_pass(const MpsCsr& csr, int64_t ridx, int lane,
                                         const float* __restrict__ wloc, float* __restrict__ gacc,
                                         float alpha, float& loss, float& correct) {
  const int64_t beg = csr.row_ptr[ridx], end = csr.row_ptr[ridx + 1];
  uint32_t c0 = 0, c1 = 0;
  float x0 = 0.f, x1 = 0.f, acc = 0.f;
  for (int64_t base = beg; base < end; base += 64) {
    const int64_t j0 = base + lane, j1 = j0 + 32;
    uint32_t kc0 = 0, kc1 = 0;
    float kx0 = 0.f, kx1 = 0.f;
    if (j0 < end) { kc0 = ld_stream(csr.col + j0); kx0 = ld_stream(csr.val + j0); }
    if (j1 < end) { kc1 = ld_stream(csr.col + j1); kx1 = ld_stream(csr.val + j1); }
    const float w0 = (j0 < end) ? ld_cg(wloc + kc0) : 0.f;
    const float w1 = (j1 < end) ? ld_cg(wloc + kc1) : 0.f;
    acc = fmaf(w0, kx0, acc);
    acc = fmaf(w1, kx1, acc);
    if (base == beg) { c0 = kc0; c1 = kc1; x0 = kx0; x1 = kx1; }
  }
  const float z = warp_sum(acc);
  const float y = csr.label[ridx];
  const float pred = sigmoidf_fast(z);
  if (lane == 0) {
    loss += logloss(z, y);
    correct += ((pred > 0.5f) == (y > 0.5f)) ? 1.f : 0.f;
  }
  if (kTrain) {
    const float g = alpha * (y - pred);
    if (beg + lane < end && x0 != 0.f) red_gpu(gacc + c0, g * x0);
    if (beg + lane + 32 < end && x1 != 0.f) red_gpu(gacc + c1, g * x1);
    for (int64_t base = beg + 64; base < end; base += 64) {
      const int64_t j0 = base + lane, j1 = j0 + 32;
      if (j0 < end) {
        const float kx = ld_stream(csr.val + j0);
        if (kx != 0.f) red_gpu(gacc + ld_stream(csr.col + j0), g * kx);
      }
      if (j1 < end) {
        const float kx = ld_stream(csr.val + j1);
        if (kx != 0.f) red_gpu(gacc + ld_stream(csr.col + j1), g * kx);
      }
    }
  }
}

__global__ void __launch_bounds__(512, 2)
lr_train_dense_kernel(const __grid_constant__ MpsTableView v, const __grid_constant__ MpsClockView c,
                      int worker, int start_clock, const __grid_constant__ MpsCsr csr,
                      const int64_t* __restrict__ batch_starts, int n_iters, int64_t batch, float alpha,
                      float* __restrict__ replica, float* __restrict__ gacc, float* __restrict__ loss_out) {
  cg::grid_group grid = cg::this_grid();
  const int64_t tid = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int64_t nthreads = static_cast<int64_t>(gridDim.x) * blockDim.x;
  const int lane = threadIdx.x & 31;
  const int64_t warp = tid >> 5, nwarps = nthreads >> 5;
  const bool bsp_owner = (c.model == MPS_MODEL_BSP) && c.own_count > 0;

  for (int it = 0; it < n_iters; ++it) {
    const int p = start_clock + it;
    // A. consistency gate for the Get of clock p (spins on the local pad only)
    cta_gate(c, p);
    // B. pull: every shard -> replica (KVClientTable::Get of all keys, fused)
    slab_pull(v, c.own_first, replica, tid, nthreads);
    grid.sync();
    // C. forward SpMV + sigmoid + gradient, accumulated locally
    float loss = 0.f, correct = 0.f;
    const int64_t row0 = batch_starts[it];
    for (int64_t r = warp; r < batch; r += nwarps) {
      int64_t ridx = row0 + r;
      if (ridx >= csr.n_rows) ridx %= csr.n_rows;
      row_pass<true>(csr, ridx, lane, replica, gacc, alpha, loss, correct);
    }
    block_accumulate(loss, correct, loss_out + 2 * it);
    grid.sync();
    // D. push: gacc -> owners (KVClientTable::Add of all touched keys, fused)
    slab_push(v, c.own_first, gacc, tid, nthreads);
    __threadfence_system();
    grid.sync();
    // E. Clock: publish progress[worker] = p + 1 on every replica
    if (tid == 0) {
      fence_sys();
      for (int i = 0; i < c.num_ranks; ++i) {
        const int r = (c.my_rank + i) % c.num_ranks;
        st_release_sys(c.pad[r] + MPS_PAD_PROGRESS + worker, p + 1);
      }
    }
    // F. BSP clock edge at the owner: wait for all workers, fold, publish
    if (bsp_owner) {
      __shared__ int ok;
      if (threadIdx.x == 0) ok = spin_until(c, [&] { return pad_min_clock(c) >= p + 1; }) ? 1 : 0;
      __syncthreads();
      if (ok) {
        for (int s = c.own_first; s < c.own_first + c.own_count; ++s) {
          if (!v.staging[s]) continue;
          const int64_t len = static_cast<int64_t>(v.begin[s + 1] - v.begin[s]);
          float* w = static_cast<float*>(v.weights[s]);
          float* st = static_cast<float*>(v.staging[s]);
          const int64_t len4 = len >> 2;
          for (int64_t i = tid; i < len4; i += nthreads) {
            const float4 d = ld_cg_v4(st + 4 * i);
            if (d.x != 0.f || d.y != 0.f || d.z != 0.f || d.w != 0.f) {
              float4 x = *reinterpret_cast<float4*>(w + 4 * i);
              x.x += d.x; x.y += d.y; x.z += d.z; x.w += d.w;
              *reinterpret_cast<float4*>(w + 4 * i) = x;
              *reinterpret_cast<float4*>(st + 4 * i) = make_float4(0.f, 0.f, 0.f, 0.f);
            }
          }
          for (int64_t i = (len4 << 2) + tid; i < len; i += nthreads) {
            const float d = ld_cg(st + i);
            if (d != 0.f) { w[i] += d; st[i] = 0.f; }
          }
        }
      }
      grid.sync();
      if (tid == 0 && ok) {
        fence_sys();
        for (int i = 0; i < c.num_ranks; ++i) {
          const int r = (c.my_rank + i) % c.num_ranks;
          for (int s = c.own_first; s < c.own_first + c.own_count; ++s)
            st_release_sys(c.pad[r] + MPS_PAD_FOLDED + s, p + 1);
        }
      }
    }
  }
}

__global__ void __launch_bounds__(256)
lr_eval_kernel(const float* __restrict__ weights, const __grid_constant__ MpsCsr csr, int64_t row_begin,
               int64_t n, float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  float loss = 0.f, correct = 0.f;
  for (int64_t r = warp; r < n; r += nwarps) {
    int64_t ridx = row_begin + r;
    if (ridx >= csr.n_rows) ridx %= csr.n_rows;
    row_pass<false>(csr, ridx, lane, weights, nullptr, 0.f, loss, correct);
  }
  block_accumulate(loss, correct, out);
}

}  // namespace mps

extern "C" {

int mps_lr_step_sparse(const MpsTableView* view, const MpsClockView* clk, int worker, int progress,
                       const MpsCsr* csr, int64_t row_begin, int64_t batch, float alpha, float* loss_out,
                       void* stream) {
  (void)worker;
  if (batch <= 0) return 0;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  int64_t want = (batch + 7) / 8;  // 8 warps (rows) per CTA
  const int grid = static_cast<int>(want < sms * 8 ? want : sms * 8);
  MpsClockView c{};
  if (clk) c = *clk;
  mps::lr_step_sparse_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      *view, c, clk != nullptr, progress, *csr, row_begin, batch, alpha, loss_out);
  MPS_LAUNCH_CHECK();
  return 0;
}

int mps_lr_train_dense(const MpsTableView* view, const MpsClockView* clk, int worker, int start_clock,
                       const MpsCsr* csr, const int64_t* batch_starts, int n_iters, int64_t batch,
                       float alpha, float* replica, float* gacc, float* loss_out, int max_ctas,
                       void* stream) {
  if (n_iters <= 0) return 0;
  if (view->dtype != MPS_F32 || view->vals_per_key != 1) return static_cast<int>(cudaErrorInvalidValue);
  int dev = 0, sms = 148, per_sm = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  MPS_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, mps::lr_train_dense_kernel, 512, 0));
  if (per_sm < 1) return static_cast<int>(cudaErrorLaunchOutOfResources);
  int grid = sms * per_sm;
  if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
  MpsTableView v = *view;
  MpsClockView c = *clk;
  MpsCsr m = *csr;
  void* args[] = {&v, &c, &worker, &start_clock, &m, &batch_starts, &n_iters, &batch, &alpha,
                  &replica, &gacc, &loss_out};
  MPS_CHECK(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(mps::lr_train_dense_kernel), dim3(grid),
                                        dim3(512), args, 0, static_cast<cudaStream_t>(stream)));
  return 0;
}

int mps_lr_eval(const float* weights, const MpsCsr* csr, int64_t row_begin, int64_t n, float* out,
                void* stream) {
  if (n <= 0) return 0;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  int64_t want = (n + 7) / 8;
  const int grid = static_cast<int>(want < sms * 8 ? want : sms * 8);
  mps::lr_eval_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(weights, *csr, row_begin, n, out);
  MPS_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"

// ------------------------------------------------------------------ allocation-free variant
// mps_lr_train_dense needs two scratch vectors (the weight replica and the local gradient
// accumulator) that its caller must allocate.  Some callers cannot: a host that has already
// handed every free byte of HBM to its own pool so a fresh cudaMalloc fails, code that runs
// where an allocation's implicit device-wide synchronisation would stall other streams
// (a latency-critical serving or control loop), or a plain C/C++ embedder that has no
// allocator of its own to hand us memory from.  For them the scratch below is part of the
// library image itself: it exists from the moment the module is loaded on a GPU until the
// process exits, on every GPU that loads it, whether or not the entry point is ever called.
#define MPS_NOALLOC_SCRATCH_FLOATS (2 * (size_t)MPS_NOALLOC_MAX_KEYS) /* replica | gacc = 8 MiB */
static __device__ __align__(256) float g_noalloc_scratch[MPS_NOALLOC_SCRATCH_FLOATS];

namespace mps {
namespace {
constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mu;
float* g_noalloc_ptr[kMaxDevices] = {nullptr};  // device address of g_noalloc_scratch, per device

// In host code `g_noalloc_scratch` is only a handle for the runtime, not an address a kernel
// can use; the real per-device address has to be asked for, with that device current.
int noalloc_scratch_for_current_device(float** out) {
  int dev = 0;
  MPS_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return static_cast<int>(cudaErrorInvalidDevice);
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (g_noalloc_ptr[dev] == nullptr) {
    float* p = nullptr;
    MPS_CHECK(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_scratch));
    if (p == nullptr) return static_cast<int>(cudaErrorInvalidSymbol);
    g_noalloc_ptr[dev] = p;
  }
  *out = g_noalloc_ptr[dev];
  return 0;
}
}  // namespace
}  // namespace mps

extern "C" int mps_lr_train_dense_noalloc(const MpsTableView* view, const MpsClockView* clk, int worker,
                                          int start_clock, const MpsCsr* csr, const int64_t* batch_starts,
                                          int n_iters, int64_t batch, float alpha, float* loss_out,
                                          int max_ctas, void* stream) {
  if (n_iters <= 0) return 0;
  if (view->dtype != MPS_F32 || view->vals_per_key != 1) return static_cast<int>(cudaErrorInvalidValue);
  const uint64_t n_keys = static_cast<uint64_t>(view->begin[view->num_shards]) - view->begin[0];
  if (n_keys > MPS_NOALLOC_MAX_KEYS) {
    fprintf(stderr,
            "mps_lr_train_dense_noalloc: table has %llu keys but the built-in scratch holds at most %llu "
            "(MPS_NOALLOC_MAX_KEYS); use mps_lr_train_dense with caller-provided scratch instead\n",
            static_cast<unsigned long long>(n_keys), static_cast<unsigned long long>(MPS_NOALLOC_MAX_KEYS));
    return static_cast<int>(cudaErrorInvalidValue);
  }
  float* scratch = nullptr;
  const int rc = mps::noalloc_scratch_for_current_device(&scratch);
  if (rc != 0) return rc;
  float* replica = scratch;
  float* gacc = scratch + MPS_NOALLOC_MAX_KEYS;
  auto st = static_cast<cudaStream_t>(stream);
  // The array still holds whatever the previous call left.  The replica is overwritten by the
  // slab pull before it is read; the gradient accumulator must start at zero.
  MPS_CHECK(cudaMemsetAsync(gacc, 0, n_keys * sizeof(float), st));
  int dev = 0, sms = 148, per_sm = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  MPS_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, mps::lr_train_dense_kernel, 512, 0));
  if (per_sm < 1) return static_cast<int>(cudaErrorLaunchOutOfResources);
  int grid = sms * per_sm;
  if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
  MpsTableView v = *view;
  MpsClockView c = *clk;
  MpsCsr m = *csr;
  void* args[] = {&v, &c, &worker, &start_clock, &m, &batch_starts, &n_iters, &batch, &alpha,
                  &replica, &gacc, &loss_out};
  MPS_CHECK(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(mps::lr_train_dense_kernel), dim3(grid),
                                        dim3(512), args, 0, st));
  return 0;
}
