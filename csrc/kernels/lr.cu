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
#include <cstdio>
#include <mutex>

#include "lr_common.cuh"

namespace mps {

// ------------------------------------------------------------------ sparse regime
// L lanes per row (32/L rows in flight per warp), up to kSparseRegs non-zeros per lane held in
// registers: all of a row's peer loads are issued before the first is consumed, and the same
// registers feed the red.add pass.
constexpr int kSparseRegs = 8;

template <int L, typename T>
__global__ void __launch_bounds__(256)
lr_step_sparse_kernel(const __grid_constant__ MpsTableView v, const __grid_constant__ MpsClockView c,
                      int gated, int progress, const __grid_constant__ MpsCsr csr, int64_t row_begin,
                      int64_t batch, T alpha, float* __restrict__ loss_out) {
  constexpr int E = kSparseRegs;
  if (gated) cta_gate(c, progress);
  const int lane = threadIdx.x & 31;
  const int gl = lane & (L - 1);
  const unsigned mask = (L == 32) ? 0xffffffffu : (((1u << L) - 1u) << (lane & ~(L - 1)));
  const int64_t group = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) / L;
  const int64_t ngroups = (static_cast<int64_t>(gridDim.x) * blockDim.x) / L;
  float loss = 0.f, correct = 0.f;
  for (int64_t r = group; r < batch; r += ngroups) {
    int64_t ridx = row_begin + r;
    if (ridx >= csr.n_rows) ridx %= csr.n_rows;
    const int64_t beg = csr.row_ptr[ridx], end = csr.row_ptr[ridx + 1];
    uint32_t off[E];
    T x[E];
    int sh[E];
#pragma unroll
    for (int e = 0; e < E; ++e) {
      const int64_t j = beg + gl + e * L;
      x[e] = T(0);
      sh[e] = -1;
      off[e] = 0;
      if (j < end) {
        const uint32_t key = ld_stream(csr.col + j);
        x[e] = static_cast<T>(ld_stream(csr.val + j));
        sh[e] = owner_of(v, key);
        if (sh[e] >= 0) off[e] = key - v.begin[sh[e]];
      }
    }
    // pull + forward (Get fused with the SpMV): peer ld straight from the owning shard
    T w[E];
#pragma unroll
    for (int e = 0; e < E; ++e) w[e] = sh[e] >= 0 ? ld_sys(static_cast<const T*>(v.weights[sh[e]]) + off[e]) : T(0);
    T acc = T(0);
#pragma unroll
    for (int e = 0; e < E; ++e) acc = fma(w[e], x[e], acc);
    for (int64_t j = beg + gl + E * L; j < end; j += L) {  // rows longer than L*E non-zeros
      const uint32_t key = ld_stream(csr.col + j);
      const int s2 = owner_of(v, key);
      if (s2 >= 0)
        acc = fma(ld_sys(static_cast<const T*>(v.weights[s2]) + (key - v.begin[s2])),
                  static_cast<T>(ld_stream(csr.val + j)), acc);
    }
    const T z = group_sum<L>(acc, mask);
    const T y = static_cast<T>(csr.label[ridx]);
    const T pred = sigmoid_t(z);
    const T g = alpha * (y - pred);
    if (gl == 0) {
      loss += static_cast<float>(logloss_t(z, y));
      correct += ((pred > T(0.5)) == (y > T(0.5))) ? 1.f : 0.f;
    }
    // gradient + push (Add fused): red.add at the owner (server/vector_storage.hpp:28-38)
#pragma unroll
    for (int e = 0; e < E; ++e)
      if (sh[e] >= 0 && x[e] != T(0))
        red_sys(static_cast<T*>(v.staging[sh[e]] ? v.staging[sh[e]] : v.weights[sh[e]]) + off[e], g * x[e]);
    for (int64_t j = beg + gl + E * L; j < end; j += L) {
      const uint32_t key = ld_stream(csr.col + j);
      const T kx = static_cast<T>(ld_stream(csr.val + j));
      const int s2 = owner_of(v, key);
      if (s2 >= 0 && kx != T(0))
        red_sys(static_cast<T*>(v.staging[s2] ? v.staging[s2] : v.weights[s2]) + (key - v.begin[s2]), g * kx);
    }
  }
  block_accumulate(loss, correct, loss_out);
}

// ------------------------------------------------------------------ dense regime
// slab-pull all shards into replica[0 .. num_dims), shard by shard (any alignment)
template <typename T>
__device__ __forceinline__ void slab_pull_by_shard(const MpsTableView& v, int rot_shard, T* __restrict__ replica,
                                                   int64_t tid, int64_t nthreads) {
  const int S = v.num_shards;
  for (int i = 0; i < S; ++i) {
    int s = rot_shard + i;
    if (s >= S) s -= S;
    const int64_t fb = flat_begin1(v, s);
    const int64_t len = static_cast<int64_t>(v.begin[s + 1] - v.begin[s]);
    const T* src = static_cast<const T*>(v.weights[s]);
    T* dst = replica + fb;
    for (int64_t q = tid; q < len; q += nthreads) dst[q] = ld_sys(src + q);
  }
}

// slab-push gacc into the owners' WEIGHTS, clearing gacc behind itself.  The fused kernel adds in
// place even under BSP: isolation comes from the "pulled" flags (every worker finished its Get of
// this clock before anyone adds), so no staging pass is needed.
template <typename T>
__device__ __forceinline__ void slab_push_by_shard(const MpsTableView& v, int rot_shard, T* __restrict__ gacc,
                                                   int64_t tid, int64_t nthreads) {
  const int S = v.num_shards;
  for (int i = 0; i < S; ++i) {
    int s = rot_shard + i;
    if (s >= S) s -= S;
    const int64_t fb = flat_begin1(v, s);
    const int64_t len = static_cast<int64_t>(v.begin[s + 1] - v.begin[s]);
    T* dst = static_cast<T*>(v.weights[s]);
    T* src = gacc + fb;
    for (int64_t q = tid; q < len; q += nthreads) {
      const T x = ld_cg(src + q);
      if (x != T(0)) { red_sys(dst + q, x); src[q] = T(0); }
    }
  }
}

// Flat versions.  The per-shard loops above hand every shard to the SAME low-numbered threads, which
// then pay one peer round trip per shard, one after the other (8 shards = 8 NVLink latencies while most
// of the grid idles).  Here the whole key range is one index space: every thread owns one or two 16-byte
// chunks, routes each to its owner and has all its loads in flight at once.  Needs every shard boundary
// on a multiple of the chunk width (4 floats / 2 doubles); otherwise the per-shard loops run.
template <typename T>
__device__ __forceinline__ void slab_pull(const MpsTableView& v, int rot_shard, T* __restrict__ replica,
                                          int64_t tid, int64_t nthreads) {
  constexpr int N = Chunk<T>::N;
  if (!shards_chunk_aligned<T>(v)) {
    slab_pull_by_shard(v, rot_shard, replica, tid, nthreads);
    return;
  }
  const int64_t total = flat_begin1(v, v.num_shards);
  const int64_t totalc = total / N;
  const int64_t rotc = flat_begin1(v, rot_shard) / N;  // start with the local shard, peers follow
  for (int64_t q = tid; q < totalc; q += 2 * nthreads) {
    int64_t g0 = q + rotc, g1 = q + nthreads + rotc;
    if (g0 >= totalc) g0 -= totalc;
    const bool two = q + nthreads < totalc;
    if (g1 >= totalc) g1 -= totalc;
    const Chunk<T> a = ld_sys_chunk(shard_elem<T>(v, g0 * N));
    Chunk<T> b = Chunk<T>::zero();
    if (two) b = ld_sys_chunk(shard_elem<T>(v, g1 * N));
    st_chunk(replica + g0 * N, a);
    if (two) st_chunk(replica + g1 * N, b);
  }
  for (int64_t f = totalc * N + tid; f < total; f += nthreads) replica[f] = ld_sys(shard_elem<T>(v, f));
}

template <typename T>
__device__ __forceinline__ void slab_push(const MpsTableView& v, int rot_shard, T* __restrict__ gacc,
                                          int64_t tid, int64_t nthreads) {
  constexpr int N = Chunk<T>::N;
  if (!shards_chunk_aligned<T>(v)) {
    slab_push_by_shard(v, rot_shard, gacc, tid, nthreads);
    return;
  }
  const int64_t total = flat_begin1(v, v.num_shards);
  const int64_t totalc = total / N;
  const int64_t rotc = flat_begin1(v, rot_shard) / N;
  for (int64_t q = tid; q < totalc; q += nthreads) {
    int64_t g = q + rotc;
    if (g >= totalc) g -= totalc;
    const Chunk<T> x = ld_cg_chunk(gacc + g * N);
    if (x.any()) {
      red_sys_chunk(shard_elem<T>(v, g * N), x);
      st_chunk(gacc + g * N, Chunk<T>::zero());
    }
  }
  for (int64_t f = totalc * N + tid; f < total; f += nthreads) {
    const T x = ld_cg(gacc + f);
    if (x != T(0)) { red_sys(shard_elem<T>(v, f), x); gacc[f] = T(0); }
  }
}

// min over the local replica of the "pulled" table
__device__ __forceinline__ int pad_min_pulled(const MpsClockView& c) {
  const int* p = c.pad[c.my_rank] + MPS_PAD_PULLED;
  int m = 0x7fffffff;
  for (int w = 0; w < c.num_workers; ++w) {
    const int q = ld_acquire_sys(p + w);
    m = q < m ? q : m;
  }
  return m;
}

// Persistent training kernel.  Per iteration (clock p):
//   A  gate:  BSP min(progress) >= p | SSP p <= min + s | ASP none      (local spins)
//   B  Get :  slab-pull all shards -> replica                        --- grid sync 1
//      publish pulled[worker] = p+1 on every rank (BSP only)
//   C  rows:  SpMV + sigmoid + gradient into gacc                    --- grid sync 2
//   D  Add :  BSP waits min(pulled) >= p+1 (every worker holds its snapshot, so in-place
//             adds cannot leak into a Get of clock p — the isolation BSPModel gets from
//             buffering adds, bsp_model.cpp:34-43), then slab-push in place
//   E  Clock: the LAST CTA to finish its part of the push publishes progress[worker] = p+1
//             on every rank, behind a system-scope fence; nobody waits for it here — the next
//             gate does.
// timeline (optional, 8 slots per iteration): globaltimer at the phase boundaries, CTA 0.
template <int L, typename T>
__global__ void __launch_bounds__(512, 2)
lr_train_dense_kernel(const __grid_constant__ MpsTableView v, const __grid_constant__ MpsClockView c,
                      int worker, int start_clock, const __grid_constant__ MpsCsr csr,
                      const int64_t* __restrict__ batch_starts, int n_iters, int64_t batch, T alpha,
                      T* __restrict__ replica, T* __restrict__ gacc, float* __restrict__ loss_out,
                      unsigned long long* __restrict__ timeline) {
  const int64_t tid = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int64_t nthreads = static_cast<int64_t>(gridDim.x) * blockDim.x;
  const bool bsp = (c.model == MPS_MODEL_BSP);
  const bool multi = c.num_ranks > 1;
  const uint32_t key0 = v.begin[0];
  const uint32_t n_keys = v.begin[v.num_shards] - key0;
  int* pad = c.pad[c.my_rank];
  GridBarrier grid{pad + MPS_PAD_LOCALCTR + 2, 0};  // zeroed by the host before the launch
  const bool tl = timeline != nullptr && tid == 0;

  for (int it = 0; it < n_iters; ++it) {
    const int p = start_clock + it;
    if (tl) timeline[8 * it + 0] = globaltimer_ns();
    // A. gate
    if (threadIdx.x == 0) {
      if (bsp) spin_until(c, [&] { return pad_min_clock(c) >= p; });
      else spin_until(c, [&] { return gate_open(c, p); });
    }
    __syncthreads();
    if (tl) timeline[8 * it + 1] = globaltimer_ns();
    // B. Get
    slab_pull(v, c.own_first, replica, tid, nthreads);
    grid.sync();
    if (tl) timeline[8 * it + 2] = globaltimer_ns();
    if (bsp && tid == 0) {
      // "my loads are done": the barrier above retired them; the fence orders them before the
      // flag for the peers whose in-place adds wait on it.
      if (multi) fence_sys();
      for (int i = 0; i < c.num_ranks; ++i) {
        const int r = (c.my_rank + i) % c.num_ranks;
        st_relaxed_sys(c.pad[r] + MPS_PAD_PULLED + worker, p + 1);
      }
    }
    // C. rows
    float loss = 0.f, correct = 0.f;
    rows_pass<L, true, T>(csr, batch_starts[it], batch, key0, n_keys, replica, gacc, alpha, false, tid / L,
                          nthreads / L, loss, correct);
    block_accumulate(loss, correct, loss_out + 2 * it);
    grid.sync();
    if (tl) timeline[8 * it + 3] = globaltimer_ns();
    // D. Add
    if (bsp) {
      if (threadIdx.x == 0) spin_until(c, [&] { return pad_min_pulled(c) >= p + 1; });
      __syncthreads();
    }
    if (tl) timeline[8 * it + 4] = globaltimer_ns();
    slab_push(v, c.own_first, gacc, tid, nthreads);
    if (tl) timeline[8 * it + 6] = globaltimer_ns();
    // E. Clock — last CTA out signals.  bar.sync orders the CTA's reds before thread 0's
    //    system fence (fence cumulativity), the atomic counter orders CTAs, and the last CTA
    //    fences at system scope again before the flag stores so that a peer's acquire load of
    //    the flag is ordered after every CTA's reds (costs one more MEMBAR.SYS per iteration,
    //    taken by a CTA whose own writes have already drained).
    __syncthreads();
    if (threadIdx.x == 0) {
      if (multi) fence_sys(); else __threadfence();
      const int done = atomicAdd(pad + MPS_PAD_LOCALCTR + 1, 1) + 1;
      if (done == static_cast<int>(gridDim.x)) {
        pad[MPS_PAD_LOCALCTR + 1] = 0;
        if (multi) fence_sys(); else __threadfence();
        for (int i = 0; i < c.num_ranks; ++i) {
          const int r = (c.my_rank + i) % c.num_ranks;
          st_relaxed_sys(c.pad[r] + MPS_PAD_PROGRESS + worker, p + 1);
        }
      }
    }
    if (tl) timeline[8 * it + 5] = globaltimer_ns();
  }

  // Keep the generic Get gate (folded[s] >= clock) usable after a fused run: once every
  // worker reached the final clock the owner republishes folded for its shards (nothing
  // is left in staging — the fused kernel never writes there).
  if (bsp && c.own_count > 0 && tid == 0) {
    const int fin = start_clock + n_iters;
    if (spin_until(c, [&] { return pad_min_clock(c) >= fin; })) {
      if (multi) fence_sys(); else __threadfence();
      for (int i = 0; i < c.num_ranks; ++i) {
        const int r = (c.my_rank + i) % c.num_ranks;
        for (int s = c.own_first; s < c.own_first + c.own_count; ++s)
          st_relaxed_sys(c.pad[r] + MPS_PAD_FOLDED + s, fin);
      }
    }
  }
}

template <int L, typename T>
__global__ void __launch_bounds__(256)
lr_eval_kernel(const T* __restrict__ weights, uint32_t key0, uint32_t n_keys, const __grid_constant__ MpsCsr csr,
               int64_t row_begin, int64_t n, float* __restrict__ out) {
  const int64_t tid = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int64_t nthreads = static_cast<int64_t>(gridDim.x) * blockDim.x;
  float loss = 0.f, correct = 0.f;
  rows_pass<L, false, T>(csr, row_begin, n, key0, n_keys, weights, nullptr, T(0), false, tid / L, nthreads / L,
                         loss, correct);
  block_accumulate(loss, correct, out);
}

template <typename T>
static void* dense_kernel_for(int L) {
  return L == 8 ? reinterpret_cast<void*>(lr_train_dense_kernel<8, T>)
                : (L == 16 ? reinterpret_cast<void*>(lr_train_dense_kernel<16, T>)
                           : reinterpret_cast<void*>(lr_train_dense_kernel<32, T>));
}

template <typename T>
static int launch_dense(const MpsTableView* view, const MpsClockView* clk, int worker, int start_clock,
                        const MpsCsr* csr, const int64_t* batch_starts, int n_iters, int64_t batch,
                        double alpha_f, void* replica_v, void* gacc_v, float* loss_out,
                        unsigned long long* timeline, int avg_nnz, int max_ctas, cudaStream_t st) {
  int dev = 0, sms = 148, per_sm = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int L = lanes_for(avg_nnz > 0 ? avg_nnz : 40);
  void* fn = dense_kernel_for<T>(L);
  MPS_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, 512, 0));
  if (per_sm < 1) return static_cast<int>(cudaErrorLaunchOutOfResources);
  int grid = sms * per_sm;
  if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
  MpsTableView v = *view;
  MpsClockView c = *clk;
  MpsCsr m = *csr;
  T alpha = static_cast<T>(alpha_f);
  T* replica = static_cast<T*>(replica_v);
  T* gacc = static_cast<T*>(gacc_v);
  void* args[] = {&v, &c, &worker, &start_clock, &m, &batch_starts, &n_iters, &batch, &alpha,
                  &replica, &gacc, &loss_out, &timeline};
  // arrival counter of the in-kernel grid barrier starts at zero
  MPS_CHECK(cudaMemsetAsync(c.pad[c.my_rank] + MPS_PAD_LOCALCTR + 2, 0, sizeof(int), st));
  MPS_CHECK(cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(512), args, 0, st));
  return 0;
}

static int launch_dense_any(const MpsTableView* view, const MpsClockView* clk, int worker, int start_clock,
                            const MpsCsr* csr, const int64_t* batch_starts, int n_iters, int64_t batch,
                            double alpha, void* replica, void* gacc, float* loss_out,
                            unsigned long long* timeline, int avg_nnz, int max_ctas, cudaStream_t st) {
  if (view->vals_per_key != 1 || csr->format != MPS_CSR_PLAIN) return static_cast<int>(cudaErrorInvalidValue);
  if (view->dtype == MPS_F64)
    return launch_dense<double>(view, clk, worker, start_clock, csr, batch_starts, n_iters, batch, alpha, replica,
                                gacc, loss_out, timeline, avg_nnz, max_ctas, st);
  if (view->dtype == MPS_F32)
    return launch_dense<float>(view, clk, worker, start_clock, csr, batch_starts, n_iters, batch, alpha, replica,
                               gacc, loss_out, timeline, avg_nnz, max_ctas, st);
  return static_cast<int>(cudaErrorInvalidValue);
}

}  // namespace mps

extern "C" {

int mps_lr_step_sparse(const MpsTableView* view, const MpsClockView* clk, int worker, int progress,
                       const MpsCsr* csr, int64_t row_begin, int64_t batch, double alpha, float* loss_out,
                       void* stream) {
  (void)worker;
  if (batch <= 0) return 0;
  if (view->vals_per_key != 1 || (view->dtype != MPS_F32 && view->dtype != MPS_F64) || csr->format != MPS_CSR_PLAIN)
    return static_cast<int>(cudaErrorInvalidValue);
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  int64_t want = (batch + 31) / 32;  // 32 row groups of 8 lanes per 256-thread CTA
  const int grid = static_cast<int>(want < 1 ? 1 : (want < sms * 8 ? want : sms * 8));
  MpsClockView c{};
  if (clk) c = *clk;
  auto st = static_cast<cudaStream_t>(stream);
  if (view->dtype == MPS_F64)
    mps::lr_step_sparse_kernel<8, double><<<grid, 256, 0, st>>>(*view, c, clk != nullptr, progress, *csr, row_begin,
                                                                batch, alpha, loss_out);
  else
    mps::lr_step_sparse_kernel<8, float><<<grid, 256, 0, st>>>(*view, c, clk != nullptr, progress, *csr, row_begin,
                                                               batch, static_cast<float>(alpha), loss_out);
  MPS_LAUNCH_CHECK();
  return 0;
}

int mps_lr_train_dense(const MpsTableView* view, const MpsClockView* clk, int worker, int start_clock,
                       const MpsCsr* csr, const int64_t* batch_starts, int n_iters, int64_t batch,
                       double alpha, void* replica, void* gacc, float* loss_out, int max_ctas,
                       void* stream) {
  if (n_iters <= 0) return 0;
  return mps::launch_dense_any(view, clk, worker, start_clock, csr, batch_starts, n_iters, batch, alpha, replica,
                               gacc, loss_out, nullptr, 0, max_ctas, static_cast<cudaStream_t>(stream));
}

int mps_lr_train_dense_ex(const MpsTableView* view, const MpsClockView* clk, int worker, int start_clock,
                          const MpsCsr* csr, const int64_t* batch_starts, int n_iters, int64_t batch,
                          double alpha, void* replica, void* gacc, float* loss_out,
                          unsigned long long* timeline, int avg_nnz, int max_ctas, void* stream) {
  if (n_iters <= 0) return 0;
  return mps::launch_dense_any(view, clk, worker, start_clock, csr, batch_starts, n_iters, batch, alpha, replica,
                               gacc, loss_out, timeline, avg_nnz, max_ctas, static_cast<cudaStream_t>(stream));
}

int mps_lr_eval(const void* weights, int dtype, uint32_t key0, uint32_t n_keys, const MpsCsr* csr,
                int64_t row_begin, int64_t n, float* out, void* stream) {
  if (n <= 0) return 0;
  if (csr->format != MPS_CSR_PLAIN) return static_cast<int>(cudaErrorInvalidValue);
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  int64_t want = (n + 31) / 32;  // 32 row groups of 8 lanes per 256-thread CTA
  const int grid = static_cast<int>(want < sms * 8 ? want : sms * 8);
  auto st = static_cast<cudaStream_t>(stream);
  if (dtype == MPS_F64)
    mps::lr_eval_kernel<8, double><<<grid, 256, 0, st>>>(static_cast<const double*>(weights), key0, n_keys, *csr,
                                                         row_begin, n, out);
  else if (dtype == MPS_F32)
    mps::lr_eval_kernel<8, float><<<grid, 256, 0, st>>>(static_cast<const float*>(weights), key0, n_keys, *csr,
                                                        row_begin, n, out);
  else
    return static_cast<int>(cudaErrorInvalidValue);
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
  return mps::launch_dense<float>(view, clk, worker, start_clock, csr, batch_starts, n_iters, batch, alpha, replica,
                                  gacc, loss_out, nullptr, 0, max_ctas, st);
}
