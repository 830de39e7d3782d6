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

template <int L>
__device__ __forceinline__ float group_sum_sparse(float v, unsigned mask) {
#pragma unroll
  for (int o = L / 2; o > 0; o >>= 1) v += __shfl_xor_sync(mask, v, o);
  return v;
}

// ------------------------------------------------------------------ sparse regime
// L lanes per row (32/L rows in flight per warp), up to kSparseRegs non-zeros per lane held in
// registers: all of a row's peer loads are issued before the first is consumed, and the same
// registers feed the red.add pass.
constexpr int kSparseRegs = 8;

template <int L>
__global__ void __launch_bounds__(256)
lr_step_sparse_kernel(const __grid_constant__ MpsTableView v, const __grid_constant__ MpsClockView c,
                      int gated, int progress, const __grid_constant__ MpsCsr csr, int64_t row_begin,
                      int64_t batch, float alpha, float* __restrict__ loss_out) {
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
    float x[E];
    int sh[E];
#pragma unroll
    for (int e = 0; e < E; ++e) {
      const int64_t j = beg + gl + e * L;
      x[e] = 0.f;
      sh[e] = -1;
      off[e] = 0;
      if (j < end) {
        const uint32_t key = ld_stream(csr.col + j);
        x[e] = ld_stream(csr.val + j);
        sh[e] = owner_of(v, key);
        if (sh[e] >= 0) off[e] = key - v.begin[sh[e]];
      }
    }
    // pull + forward (Get fused with the SpMV): peer ld straight from the owning shard
    float w[E];
#pragma unroll
    for (int e = 0; e < E; ++e) w[e] = sh[e] >= 0 ? ld_sys(static_cast<const float*>(v.weights[sh[e]]) + off[e]) : 0.f;
    float acc = 0.f;
#pragma unroll
    for (int e = 0; e < E; ++e) acc = fmaf(w[e], x[e], acc);
    for (int64_t j = beg + gl + E * L; j < end; j += L) {  // rows longer than L*E non-zeros
      const uint32_t key = ld_stream(csr.col + j);
      const int s2 = owner_of(v, key);
      if (s2 >= 0) acc = fmaf(ld_sys(static_cast<const float*>(v.weights[s2]) + (key - v.begin[s2])), ld_stream(csr.val + j), acc);
    }
    const float z = group_sum_sparse<L>(acc, mask);
    const float y = csr.label[ridx];
    const float pred = sigmoidf_fast(z);
    const float g = alpha * (y - pred);
    if (gl == 0) {
      loss += logloss(z, y);
      correct += ((pred > 0.5f) == (y > 0.5f)) ? 1.f : 0.f;
    }
    // gradient + push (Add fused): red.add at the owner (server/vector_storage.hpp:28-38)
#pragma unroll
    for (int e = 0; e < E; ++e)
      if (sh[e] >= 0 && x[e] != 0.f)
        red_sys(static_cast<float*>(v.staging[sh[e]] ? v.staging[sh[e]] : v.weights[sh[e]]) + off[e], g * x[e]);
    for (int64_t j = beg + gl + E * L; j < end; j += L) {
      const uint32_t key = ld_stream(csr.col + j);
      const float kx = ld_stream(csr.val + j);
      const int s2 = owner_of(v, key);
      if (s2 >= 0 && kx != 0.f)
        red_sys(static_cast<float*>(v.staging[s2] ? v.staging[s2] : v.weights[s2]) + (key - v.begin[s2]), g * kx);
    }
  }
  block_accumulate(loss, correct, loss_out);
}

// ------------------------------------------------------------------ dense regime
__device__ __forceinline__ int64_t flat_begin1(const MpsTableView& v, int s) {
  return static_cast<int64_t>(v.begin[s] - v.begin[0]);
}

// slab-pull all shards into replica[0 .. num_dims)
__device__ __forceinline__ void slab_pull_by_shard(const MpsTableView& v, int rot_shard, float* __restrict__ replica,
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

// slab-push gacc into the owners' WEIGHTS (vector reds), clearing gacc behind itself.  The
// fused kernel adds in place even under BSP: isolation comes from the "pulled" flags (every
// worker finished its Get of this clock before anyone adds), so no staging pass is needed.
__device__ __forceinline__ void slab_push_by_shard(const MpsTableView& v, int rot_shard, float* __restrict__ gacc,
                                                   int64_t tid, int64_t nthreads) {
  const int S = v.num_shards;
  for (int i = 0; i < S; ++i) {
    int s = rot_shard + i;
    if (s >= S) s -= S;
    const int64_t fb = flat_begin1(v, s);
    const int64_t len = static_cast<int64_t>(v.begin[s + 1] - v.begin[s]);
    float* dst = static_cast<float*>(v.weights[s]);
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

// Flat versions.  The per-shard loops above hand every shard to the SAME low-numbered threads, which
// then pay one peer round trip per shard, one after the other (8 shards = 8 NVLink latencies while most
// of the grid idles).  Here the whole key range is one index space: every thread owns one or two 16-byte
// chunks, routes each to its owner and has all its loads in flight at once.  Needs every shard boundary
// on a multiple of 4 values; otherwise the per-shard loops run.
__device__ __forceinline__ bool shards_aligned4(const MpsTableView& v) {
  bool ok = true;
  for (int s = 1; s < v.num_shards; ++s) ok = ok && ((flat_begin1(v, s) & 3) == 0);
  return ok;
}

__device__ __forceinline__ const float* chunk_src(const MpsTableView& v, int64_t f) {
  const int s = owner_of(v, v.begin[0] + static_cast<uint32_t>(f));
  return static_cast<const float*>(v.weights[s]) + (f - flat_begin1(v, s));
}

__device__ __forceinline__ void slab_pull(const MpsTableView& v, int rot_shard, float* __restrict__ replica,
                                          int64_t tid, int64_t nthreads) {
  if (!shards_aligned4(v)) {
    slab_pull_by_shard(v, rot_shard, replica, tid, nthreads);
    return;
  }
  const int64_t total = flat_begin1(v, v.num_shards);
  const int64_t total4 = total >> 2;
  const int64_t rot4 = flat_begin1(v, rot_shard) >> 2;  // start with the local shard, peers follow
  for (int64_t q = tid; q < total4; q += 2 * nthreads) {
    int64_t g0 = q + rot4, g1 = q + nthreads + rot4;
    if (g0 >= total4) g0 -= total4;
    const bool two = q + nthreads < total4;
    if (g1 >= total4) g1 -= total4;
    const float4 a = ld_sys_v4(chunk_src(v, g0 << 2));
    float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
    if (two) b = ld_sys_v4(chunk_src(v, g1 << 2));
    *reinterpret_cast<float4*>(replica + (g0 << 2)) = a;
    if (two) *reinterpret_cast<float4*>(replica + (g1 << 2)) = b;
  }
  for (int64_t f = (total4 << 2) + tid; f < total; f += nthreads) replica[f] = ld_sys(chunk_src(v, f));
}

__device__ __forceinline__ void slab_push(const MpsTableView& v, int rot_shard, float* __restrict__ gacc,
                                          int64_t tid, int64_t nthreads) {
  if (!shards_aligned4(v)) {
    slab_push_by_shard(v, rot_shard, gacc, tid, nthreads);
    return;
  }
  const int64_t total = flat_begin1(v, v.num_shards);
  const int64_t total4 = total >> 2;
  const int64_t rot4 = flat_begin1(v, rot_shard) >> 2;
  for (int64_t q = tid; q < total4; q += nthreads) {
    int64_t g = q + rot4;
    if (g >= total4) g -= total4;
    const float4 x = ld_cg_v4(gacc + (g << 2));
    if (x.x != 0.f || x.y != 0.f || x.z != 0.f || x.w != 0.f) {
      red_sys_v4(const_cast<float*>(chunk_src(v, g << 2)), x);
      *reinterpret_cast<float4*>(gacc + (g << 2)) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
  for (int64_t f = (total4 << 2) + tid; f < total; f += nthreads) {
    const float x = ld_cg(gacc + f);
    if (x != 0.f) { red_sys(const_cast<float*>(chunk_src(v, f)), x); gacc[f] = 0.f; }
  }
}

// Rows of a batch against a local dense copy of the weights.  L lanes cooperate on one row
// (32/L rows in flight per warp), each lane keeps up to E of the row's non-zeros in
// registers so the weight gathers of a row are all issued before the first one is used.
// kTrain: also scatter alpha*x*(y-pred) into gacc (device-scope red into the local L2).
constexpr int kRowRegs = 8;

template <int L>
__device__ __forceinline__ float group_sum(float v, unsigned mask) {
#pragma unroll
  for (int o = L / 2; o > 0; o >>= 1) v += __shfl_xor_sync(mask, v, o);
  return v;
}

template <int L, bool kTrain>
__device__ __forceinline__ void rows_pass(const MpsCsr& csr, int64_t row0, int64_t batch,
                                          const float* __restrict__ wloc, float* __restrict__ gacc,
                                          float alpha, int64_t group, int64_t ngroups, float& loss,
                                          float& correct) {
  constexpr int E = kRowRegs;
  const int lane = threadIdx.x & 31;
  const int gl = lane & (L - 1);
  const unsigned mask = (L == 32) ? 0xffffffffu : (((1u << L) - 1u) << (lane & ~(L - 1)));
  for (int64_t r = group; r < batch; r += ngroups) {
    int64_t ridx = row0 + r;
    if (ridx >= csr.n_rows) ridx %= csr.n_rows;
    const int64_t beg = csr.row_ptr[ridx], end = csr.row_ptr[ridx + 1];
    uint32_t c[E];
    float x[E];
#pragma unroll
    for (int e = 0; e < E; ++e) {
      const int64_t j = beg + gl + e * L;
      c[e] = 0;
      x[e] = 0.f;
      if (j < end) { c[e] = ld_stream(csr.col + j); x[e] = ld_stream(csr.val + j); }
    }
    float w[E];
#pragma unroll
    for (int e = 0; e < E; ++e) w[e] = (beg + gl + e * L < end) ? ld_cg(wloc + c[e]) : 0.f;
    float acc = 0.f;
#pragma unroll
    for (int e = 0; e < E; ++e) acc = fmaf(w[e], x[e], acc);
    for (int64_t j = beg + gl + E * L; j < end; j += L)  // rows longer than L*E non-zeros
      acc = fmaf(ld_cg(wloc + ld_stream(csr.col + j)), ld_stream(csr.val + j), acc);
    const float z = group_sum<L>(acc, mask);
    const float y = csr.label[ridx];
    const float pred = sigmoidf_fast(z);
    if (gl == 0) {
      loss += logloss(z, y);
      correct += ((pred > 0.5f) == (y > 0.5f)) ? 1.f : 0.f;
    }
    if (kTrain) {
      const float g = alpha * (y - pred);
#pragma unroll
      for (int e = 0; e < E; ++e)
        if (x[e] != 0.f) red_gpu(gacc + c[e], g * x[e]);
      for (int64_t j = beg + gl + E * L; j < end; j += L) {
        const float kx = ld_stream(csr.val + j);
        if (kx != 0.f) red_gpu(gacc + ld_stream(csr.col + j), g * kx);
      }
    }
  }
}

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
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
//             (release.sys) on every rank; nobody waits for it here — the next gate does.
// timeline (optional, 8 slots per iteration): globaltimer at the phase boundaries, CTA 0.
template <int L>
__global__ void __launch_bounds__(512, 2)
lr_train_dense_kernel(const __grid_constant__ MpsTableView v, const __grid_constant__ MpsClockView c,
                      int worker, int start_clock, const __grid_constant__ MpsCsr csr,
                      const int64_t* __restrict__ batch_starts, int n_iters, int64_t batch, float alpha,
                      float* __restrict__ replica, float* __restrict__ gacc, float* __restrict__ loss_out,
                      unsigned long long* __restrict__ timeline) {
  const int64_t tid = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int64_t nthreads = static_cast<int64_t>(gridDim.x) * blockDim.x;
  const bool bsp = (c.model == MPS_MODEL_BSP);
  const bool multi = c.num_ranks > 1;
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
      // nothing to publish but "my loads are done" (the barrier above retired them): relaxed
      for (int i = 0; i < c.num_ranks; ++i) {
        const int r = (c.my_rank + i) % c.num_ranks;
        st_relaxed_sys(c.pad[r] + MPS_PAD_PULLED + worker, p + 1);
      }
    }
    // C. rows
    float loss = 0.f, correct = 0.f;
    rows_pass<L, true>(csr, batch_starts[it], batch, replica, gacc, alpha, tid / L, nthreads / L, loss,
                       correct);
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
    //    system fence (fence cumulativity), the atomic counter orders CTAs.
    __syncthreads();
    if (threadIdx.x == 0) {
      // MEMBAR.SYS costs ~4 us on B200; with a single rank every observer is on this GPU
      // and a device-scope fence orders the reds just as well.
      if (multi) fence_sys(); else __threadfence();
      const int done = atomicAdd(pad + MPS_PAD_LOCALCTR + 1, 1) + 1;
      if (done == static_cast<int>(gridDim.x)) {
        pad[MPS_PAD_LOCALCTR + 1] = 0;
        // Every CTA's reds were performed system-wide (its own fence above) before its arrival
        // became visible, so a device-scope fence is enough to order the flag stores after
        // all of them; relaxed stores avoid a MEMBAR.SYS per store.
        __threadfence();
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

template <int L>
__global__ void __launch_bounds__(256)
lr_eval_kernel(const float* __restrict__ weights, const __grid_constant__ MpsCsr csr, int64_t row_begin,
               int64_t n, float* __restrict__ out) {
  const int64_t tid = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int64_t nthreads = static_cast<int64_t>(gridDim.x) * blockDim.x;
  float loss = 0.f, correct = 0.f;
  rows_pass<L, false>(csr, row_begin, n, weights, nullptr, 0.f, tid / L, nthreads / L, loss, correct);
  block_accumulate(loss, correct, out);
}

// lanes per row from the average row length: 8 lanes hold rows of <= 64 nnz in registers
static int lanes_for(int64_t avg_nnz) {
  if (avg_nnz <= 8 * kRowRegs) return 8;
  if (avg_nnz <= 16 * kRowRegs) return 16;
  return 32;
}

static int launch_dense(const MpsTableView* view, const MpsClockView* clk, int worker, int start_clock,
                        const MpsCsr* csr, const int64_t* batch_starts, int n_iters, int64_t batch,
                        float alpha, float* replica, float* gacc, float* loss_out,
                        unsigned long long* timeline, int avg_nnz, int max_ctas, cudaStream_t st) {
  int dev = 0, sms = 148, per_sm = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int L = lanes_for(avg_nnz > 0 ? avg_nnz : 40);
  void* fn = L == 8 ? reinterpret_cast<void*>(lr_train_dense_kernel<8>)
                    : (L == 16 ? reinterpret_cast<void*>(lr_train_dense_kernel<16>)
                               : reinterpret_cast<void*>(lr_train_dense_kernel<32>));
  MPS_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, 512, 0));
  if (per_sm < 1) return static_cast<int>(cudaErrorLaunchOutOfResources);
  int grid = sms * per_sm;
  if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
  MpsTableView v = *view;
  MpsClockView c = *clk;
  MpsCsr m = *csr;
  void* args[] = {&v, &c, &worker, &start_clock, &m, &batch_starts, &n_iters, &batch, &alpha,
                  &replica, &gacc, &loss_out, &timeline};
  // arrival counter of the in-kernel grid barrier starts at zero
  MPS_CHECK(cudaMemsetAsync(c.pad[c.my_rank] + MPS_PAD_LOCALCTR + 2, 0, sizeof(int), st));
  MPS_CHECK(cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(512), args, 0, st));
  return 0;
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
  int64_t want = (batch + 31) / 32;  // 32 row groups of 8 lanes per 256-thread CTA
  const int grid = static_cast<int>(want < 1 ? 1 : (want < sms * 8 ? want : sms * 8));
  MpsClockView c{};
  if (clk) c = *clk;
  mps::lr_step_sparse_kernel<8><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
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
  return mps::launch_dense(view, clk, worker, start_clock, csr, batch_starts, n_iters, batch, alpha, replica,
                           gacc, loss_out, nullptr, 0, max_ctas, static_cast<cudaStream_t>(stream));
}

int mps_lr_train_dense_ex(const MpsTableView* view, const MpsClockView* clk, int worker, int start_clock,
                          const MpsCsr* csr, const int64_t* batch_starts, int n_iters, int64_t batch,
                          float alpha, float* replica, float* gacc, float* loss_out,
                          unsigned long long* timeline, int avg_nnz, int max_ctas, void* stream) {
  if (n_iters <= 0) return 0;
  if (view->dtype != MPS_F32 || view->vals_per_key != 1) return static_cast<int>(cudaErrorInvalidValue);
  return mps::launch_dense(view, clk, worker, start_clock, csr, batch_starts, n_iters, batch, alpha, replica,
                           gacc, loss_out, timeline, avg_nnz, max_ctas, static_cast<cudaStream_t>(stream));
}

int mps_lr_eval(const float* weights, const MpsCsr* csr, int64_t row_begin, int64_t n, float* out,
                void* stream) {
  if (n <= 0) return 0;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  int64_t want = (n + 31) / 32;  // 32 row groups of 8 lanes per 256-thread CTA
  const int grid = static_cast<int>(want < sms * 8 ? want : sms * 8);
  mps::lr_eval_kernel<8><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(weights, *csr, row_begin, n, out);
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
  return mps::launch_dense(view, clk, worker, start_clock, csr, batch_starts, n_iters, batch, alpha, replica,
                           gacc, loss_out, nullptr, 0, max_ctas, st);
}
