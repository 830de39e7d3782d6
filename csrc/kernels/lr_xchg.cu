// lr_xchg.cu — BSP sparse-LR iterations whose Get / Add / Clock are ONE exchange per step:
// reduce at the owner, then the owner broadcasts.
//
// Reference semantics (server/consistency/bsp_model.cpp:14-43): adds of clock p are buffered and
// applied only when every worker has clocked; a Get of clock p+1 is answered after that.  The
// generic device path does it as pull -> compute -> push -> clock with three cross-GPU waits per
// iteration and the whole table crossing NVLink twice (lr.cu).  Here every rank keeps, in a
// symmetric buffer that every GPU maps (and, with NVLS, the switch maps as one multicast address),
//   replica   a full copy of the weights, kept current by the shard owners,
//   gacc[3]   its own gradient of the running batch (triple buffered).
// Iteration `it` (clock p) on every rank (G = CTAs per rank, the same on every rank):
//   A  wait    arrived[o] >= it * G for every owner o: all CTAs of every owner have finished the
//              broadcast of clock p (local spin; nothing to wait for in the first iteration)
//      zero gacc[(p+1)&1]  (its last readers, the owner phases of clock p-1, are the ones just
//                           waited for)
//   B  rows    SpMV + sigmoid + gradient of the batch -> gacc[p&1]
//   C  Clock   grid barrier, one fence.sys, progress[worker] = p+1 on every rank
//   D  owner   wait min(progress) >= p+1; for every 16-byte chunk of MY keys:
//                 sum = multimem.ld_reduce(gacc[p&1])   (the switch adds the N GPUs' copies)
//                 w   = weights[chunk] + sum -> weights (the shard itself, master copy)
//                 multimem.st(replica[chunk], w)        (lands in all N replicas)
//              without NVLS (or where it is slower): N-1 peer loads / N-1 peer stores per chunk.
//   E  signal  every CTA: bar.sync, fence.sys, arrived[me] += 1 on every rank.
// The broadcast has no "last CTA" hop: each CTA publishes its own completion behind its own system
// fence, all CTAs in parallel, and a reader that sees the full count has synchronised with every
// CTA's fence through the chain of reds on the counter.  (fence.acq_rel.sys costs 3.3 us on
// B200 even with nothing outstanding — profiles/nvlink_probe_2gpu.txt — so the protocol is built
// to have exactly two of them on an iteration's critical path.)
// After the last iteration every rank has seen every owner's last broadcast and marks, in its OWN
// pad, every worker clocked and every shard folded at the final clock, so the generic tracker
// (Get gate, checkpoint) sees the run.
// The counters live in two sets used by alternate launches; a launch clears the other set for its
// successor (safe: every signal of the previous launch was consumed before that launch ended).
// Per iteration and GPU with NVLS: |table|/N bytes in (reduce), |table|/N out (broadcast), plus
// the (N-1)/N |table| that every replica has to receive.
//
// Requires: BSP, every rank of the table runs exactly one worker through this kernel with the
// same n_iters and the same grid size, one value per key.
#include <mutex>

#include "lr_common.cuh"

namespace mps {

// ---------------------------------------------------------------- owner phase
template <typename T>
__device__ __forceinline__ void bcast_chunk(const MpsClockView& c, const MpsXchg& x, T* rmc, int64_t f,
                                            const Chunk<T>& w) {
  if (rmc) {
    if constexpr (sizeof(T) == 4) {
      mc_st_chunk(rmc + f, w);
    } else {  // one 16-byte multicast store (multimem.st has no v2.f64; the payload is just bits)
      asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(rmc + f),
                   "f"(__int_as_float(__double2loint(w.v.x))), "f"(__int_as_float(__double2hiint(w.v.x))),
                   "f"(__int_as_float(__double2loint(w.v.y))), "f"(__int_as_float(__double2hiint(w.v.y)))
                   : "memory");
    }
    return;
  }
  st_chunk(static_cast<T*>(x.replica[c.my_rank]) + f, w);
  for (int k = 1; k < c.num_ranks; ++k) {
    int r = c.my_rank + k;
    if (r >= c.num_ranks) r -= c.num_ranks;
    st_sys_chunk(static_cast<T*>(x.replica[r]) + f, w);
  }
}

template <typename T>
__device__ __forceinline__ void bcast_elem(const MpsClockView& c, const MpsXchg& x, T* rmc, int64_t f, T w) {
  if (rmc) { mc_st(rmc + f, w); return; }
  static_cast<T*>(x.replica[c.my_rank])[f] = w;
  for (int k = 1; k < c.num_ranks; ++k) {
    int r = c.my_rank + k;
    if (r >= c.num_ranks) r -= c.num_ranks;
    st_sys(static_cast<T*>(x.replica[r]) + f, w);
  }
}

// flat elements [ob, oe) of the table: reduce the N accumulators, apply, broadcast
template <typename T>
__device__ __forceinline__ void owner_reduce_bcast(const MpsTableView& v, const MpsClockView& c, const MpsXchg& x,
                                                   int buf, int64_t ob, int64_t oe, int64_t tid, int64_t nthreads) {
  constexpr int N = Chunk<T>::N;
  const int R = c.num_ranks;
  const int64_t goff = static_cast<int64_t>(buf) * x.gacc_stride;
  const T* gl = static_cast<const T*>(x.gacc[c.my_rank]) + goff;
  const T* gmc =
      (x.gacc_mc && (x.flags & MPS_XCHG_REDUCE_P2P) == 0) ? static_cast<const T*>(x.gacc_mc) + goff : nullptr;
  T* rmc = (x.replica_mc && (x.flags & MPS_XCHG_BCAST_P2P) == 0) ? static_cast<T*>(x.replica_mc) : nullptr;
  if (shards_chunk_aligned<T>(v) && (ob % N) == 0) {
    const int64_t nc = (oe - ob) / N;
    // two chunks per thread and round: all of a thread's loads are in flight before the first store
    for (int64_t i = tid; i < nc; i += 2 * nthreads) {
      const int64_t f0 = ob + i * N, f1 = f0 + nthreads * N;
      const bool two = i + nthreads < nc;
      Chunk<T> s0, s1 = Chunk<T>::zero();
      if (gmc) {
        s0 = mc_ld_reduce_chunk(gmc + f0);
        if (two) s1 = mc_ld_reduce_chunk(gmc + f1);
      } else {
        s0 = ld_cg_chunk(gl + f0);
        if (two) s1 = ld_cg_chunk(gl + f1);
        for (int k = 1; k < R; ++k) {
          int r = c.my_rank + k;
          if (r >= R) r -= R;
          const T* gp = static_cast<const T*>(x.gacc[r]) + goff;
          const Chunk<T> a = ld_sys_chunk(gp + f0);
          Chunk<T> b = Chunk<T>::zero();
          if (two) b = ld_sys_chunk(gp + f1);
          s0.add(a);
          s1.add(b);
        }
      }
      T* wp0 = shard_elem<T>(v, f0);
      Chunk<T> w0;
      w0.v = *reinterpret_cast<const decltype(w0.v)*>(wp0);
      w0.add(s0);
      st_chunk(wp0, w0);
      bcast_chunk<T>(c, x, rmc, f0, w0);
      if (two) {
        T* wp1 = shard_elem<T>(v, f1);
        Chunk<T> w1;
        w1.v = *reinterpret_cast<const decltype(w1.v)*>(wp1);
        w1.add(s1);
        st_chunk(wp1, w1);
        bcast_chunk<T>(c, x, rmc, f1, w1);
      }
    }
    ob += nc * N;  // tail below
  }
  for (int64_t f = ob + tid; f < oe; f += nthreads) {
    T sum;
    if (gmc) {
      sum = mc_ld_reduce(gmc + f);
    } else {
      sum = ld_cg(gl + f);
      for (int k = 1; k < R; ++k) {
        int r = c.my_rank + k;
        if (r >= R) r -= R;
        sum += ld_sys(static_cast<const T*>(x.gacc[r]) + goff + f);
      }
    }
    T* wp = shard_elem<T>(v, f);
    const T w = *wp + sum;
    *wp = w;
    bcast_elem<T>(c, x, rmc, f, w);
  }
}

template <typename T>
__device__ __forceinline__ void zero_range(T* p, int64_t n, int64_t tid, int64_t nthreads) {
  constexpr int N = Chunk<T>::N;
  const int64_t nc = n / N;
  for (int64_t i = tid; i < nc; i += nthreads) st_chunk(p + i * N, Chunk<T>::zero());
  for (int64_t f = nc * N + tid; f < n; f += nthreads) p[f] = T(0);
}

// the slab pull of lr.cu, used once per launch when the replica is not known to be current
template <typename T>
__device__ __forceinline__ void replica_fill(const MpsTableView& v, int rot_shard, T* __restrict__ replica,
                                             int64_t tid, int64_t nthreads) {
  constexpr int N = Chunk<T>::N;
  const int64_t total = flat_begin1(v, v.num_shards);
  if (shards_chunk_aligned<T>(v)) {
    const int64_t totalc = total / N;
    const int64_t rotc = flat_begin1(v, rot_shard) / N;
    for (int64_t i = tid; i < totalc; i += nthreads) {
      int64_t g = i + rotc;
      if (g >= totalc) g -= totalc;
      st_chunk(replica + g * N, ld_sys_chunk(shard_elem<T>(v, g * N)));
    }
    for (int64_t f = totalc * N + tid; f < total; f += nthreads) replica[f] = ld_sys(shard_elem<T>(v, f));
  } else {
    for (int64_t f = tid; f < total; f += nthreads) replica[f] = ld_sys(shard_elem<T>(v, f));
  }
}

template <int L, typename T, bool kPacked>
__global__ void __launch_bounds__(512, 2)
lr_train_xchg_kernel(const __grid_constant__ MpsTableView v, const __grid_constant__ MpsClockView c,
                     int worker, int start_clock, const __grid_constant__ MpsCsr csr,
                     const int64_t* __restrict__ batch_starts, int n_iters, int64_t batch, T alpha,
                     const __grid_constant__ MpsXchg x, float* __restrict__ loss_out,
                     unsigned long long* __restrict__ timeline) {
  const int64_t tid = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int64_t nthreads = static_cast<int64_t>(gridDim.x) * blockDim.x;
  const bool multi = c.num_ranks > 1;
  const uint32_t key0 = v.begin[0];
  const uint32_t n_keys = v.begin[v.num_shards] - key0;
  const int G = static_cast<int>(gridDim.x);
  int* pad = c.pad[c.my_rank];
  const int set = MPS_PAD_XCNT + (x.seq & 1) * 2 * MPS_MAX_RANKS;
  const int kReady = set, kArrived = set + MPS_MAX_RANKS;
  const bool tl = timeline != nullptr && tid == 0;
  T* replica = static_cast<T*>(x.replica[c.my_rank]);
  T* gacc = static_cast<T*>(x.gacc[c.my_rank]);
  const int64_t ob = c.own_count > 0 ? flat_begin1(v, c.own_first) : 0;
  const int64_t oe = c.own_count > 0 ? flat_begin1(v, c.own_first + c.own_count) : 0;

  GridBarrier grid{pad + MPS_PAD_LOCALCTR + 2, 0};  // zeroed by the host before the launch
  if (!x.replica_valid) {
    // First launch on this table (or the weights were changed behind the replica's back): wait
    // for the Get gate of start_clock, copy every shard once, start from a clean accumulator.
    if (threadIdx.x == 0) spin_until(c, [&] { return gate_open(c, start_clock); });
    __syncthreads();
    replica_fill<T>(v, c.own_first, replica, tid, nthreads);
    zero_range(gacc + static_cast<int64_t>(start_clock & 1) * x.gacc_stride, static_cast<int64_t>(n_keys), tid,
               nthreads);
    grid.sync();
  }

  for (int it = 0; it <= n_iters; ++it) {
    const int p = start_clock + it;
    const int buf = p & 1;
    if (tl && it < n_iters) timeline[8 * it + 0] = globaltimer_ns();
    // A. every owner's broadcast of this clock has landed.  A timed-out wait raises the pad error
    //    flag and every later wait falls through, so a lost peer ends the launch with an error on
    //    the host instead of a hung GPU.
    if (it > 0) {
      if (threadIdx.x == 0) spin_until(c, [&] { return all_reached(pad + kArrived, c.num_ranks, it * G); });
      __syncthreads();
    }
    if (it == n_iters) break;
    if (tl) timeline[8 * it + 1] = globaltimer_ns();
    zero_range(gacc + static_cast<int64_t>(buf ^ 1) * x.gacc_stride, static_cast<int64_t>(n_keys), tid, nthreads);
    // B. rows
    float loss = 0.f, correct = 0.f;
    rows_pass<L, true, T, kPacked>(csr, batch_starts[it], batch, key0, n_keys, replica,
                          gacc + static_cast<int64_t>(buf) * x.gacc_stride, alpha, multi, tid / L, nthreads / L, loss,
                          correct);
    block_accumulate(loss, correct, loss_out + 2 * it);  // ends with bar.sync: the CTA's reds are issued
    if (tl) timeline[8 * it + 2] = globaltimer_ns();
    // C + D. Clock, then wait for every rank's gradient.  Across GPUs: one grid barrier, ONE system
    //        fence by one thread, progress[worker] = p+1 on every rank (per-CTA fences behind the
    //        rows' reds measured 4 us slower here).  On a single GPU a device-scope fence is all a
    //        CTA needs, so every CTA counts itself in and the grid barrier is skipped.
    if (multi) {
      grid.sync();
      if (tid == 0) {
        fence_sys();
        for (int i = 0; i < c.num_ranks; ++i) {
          const int r = (c.my_rank + i) % c.num_ranks;
          st_relaxed_sys(c.pad[r] + MPS_PAD_PROGRESS + worker, p + 1);
        }
      }
      if (threadIdx.x == 0) spin_until(c, [&] { return pad_min_clock(c) >= p + 1; });
    } else if (threadIdx.x == 0) {
      cta_signal(c, kReady, false);
      spin_until(c, [&] { return all_reached(pad + kReady, 1, (it + 1) * G); });
    }
    __syncthreads();
    if (tl) timeline[8 * it + 3] = globaltimer_ns();
    if (oe > ob) owner_reduce_bcast<T>(v, c, x, buf, ob, oe, tid, nthreads);
    if (tl) timeline[8 * it + 4] = globaltimer_ns();
    // E. signal: this CTA's part of the broadcast (and of the shard update) is complete.
    __syncthreads();
    if (threadIdx.x == 0) cta_signal(c, kArrived, multi);
    if (tl) timeline[8 * it + 5] = globaltimer_ns();
  }

  // Every rank's last gradient and every owner's last broadcast / shard update have been observed
  // above: record that in my own pad for the generic tracker.
  if (tid == 0) {
    const int fin = start_clock + n_iters;
    for (int w = 0; w < c.num_workers; ++w) st_relaxed_sys(pad + MPS_PAD_PROGRESS + w, fin);
    for (int s = 0; s < c.num_shards; ++s) st_relaxed_sys(pad + MPS_PAD_FOLDED + s, fin);
  }
}

template <typename T, bool kPacked>
static void* xchg_kernel_for(int L) {
  return L == 8 ? reinterpret_cast<void*>(lr_train_xchg_kernel<8, T, kPacked>)
                : (L == 16 ? reinterpret_cast<void*>(lr_train_xchg_kernel<16, T, kPacked>)
                           : reinterpret_cast<void*>(lr_train_xchg_kernel<32, T, kPacked>));
}

template <typename T>
static int launch_xchg(const MpsTableView* view, const MpsClockView* clk, int worker, int start_clock,
                       const MpsCsr* csr, const int64_t* batch_starts, int n_iters, int64_t batch, double alpha_f,
                       const MpsXchg* xchg, float* loss_out, unsigned long long* timeline, int avg_nnz,
                       int max_ctas, cudaStream_t st) {
  const int L = lanes_for(avg_nnz > 0 ? avg_nnz : 40);
  void* fn = csr->format == MPS_CSR_PACKED ? xchg_kernel_for<T, true>(L) : xchg_kernel_for<T, false>(L);
  // the co-resident grid of a kernel variant never changes: asked once per variant and process (every launch
  // is inside somebody's timed region, and the two runtime queries cost more host time than the launch)
  static void* cached_fn[12];
  static int cached_grid[12];
  static std::mutex cache_mu;
  int grid = 0;
  {
    std::lock_guard<std::mutex> lk(cache_mu);
    int slot = 0;
    while (slot < 12 && cached_fn[slot] != nullptr && cached_fn[slot] != fn) ++slot;
    if (slot < 12 && cached_fn[slot] == fn) {
      grid = cached_grid[slot];
    } else {
      int dev = 0, sms = 148, per_sm = 0;
      cudaGetDevice(&dev);
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
      MPS_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, 512, 0));
      if (per_sm < 1) return static_cast<int>(cudaErrorLaunchOutOfResources);
      grid = sms * per_sm;
      if (slot < 12) {
        cached_fn[slot] = fn;
        cached_grid[slot] = grid;
      }
    }
  }
  if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
  MpsTableView v = *view;
  MpsClockView c = *clk;
  MpsCsr m = *csr;
  MpsXchg x = *xchg;
  T alpha = static_cast<T>(alpha_f);
  void* args[] = {&v, &c, &worker, &start_clock, &m, &batch_starts, &n_iters, &batch, &alpha, &x, &loss_out,
                  &timeline};
  // the grid barrier's counter (first-launch path) starts at zero; the counter set of the NEXT launch is
  // cleared now — all its signals from two launches ago were consumed before that launch ended, and none
  // of the next launch can be sent before this one has finished on every rank
  MPS_CHECK(cudaMemsetAsync(c.pad[c.my_rank] + MPS_PAD_LOCALCTR + 2, 0, sizeof(int), st));
  MPS_CHECK(cudaMemsetAsync(c.pad[c.my_rank] + MPS_PAD_XCNT + ((x.seq + 1) & 1) * 2 * MPS_MAX_RANKS, 0,
                            2 * MPS_MAX_RANKS * sizeof(int), st));
  MPS_CHECK(cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(512), args, 0, st));
  return 0;
}

}  // namespace mps

extern "C" int mps_lr_train_xchg(const MpsTableView* view, const MpsClockView* clk, int worker, int start_clock,
                                 const MpsCsr* csr, const int64_t* batch_starts, int n_iters, int64_t batch,
                                 double alpha, const MpsXchg* xchg, float* loss_out, unsigned long long* timeline,
                                 int avg_nnz, int max_ctas, void* stream) {
  if (n_iters <= 0) return 0;
  if (view->vals_per_key != 1 || clk->model != MPS_MODEL_BSP) return static_cast<int>(cudaErrorInvalidValue);
  auto st = static_cast<cudaStream_t>(stream);
  if (view->dtype == MPS_F64)
    return mps::launch_xchg<double>(view, clk, worker, start_clock, csr, batch_starts, n_iters, batch, alpha, xchg,
                                    loss_out, timeline, avg_nnz, max_ctas, st);
  if (view->dtype == MPS_F32)
    return mps::launch_xchg<float>(view, clk, worker, start_clock, csr, batch_starts, n_iters, batch, alpha, xchg,
                                   loss_out, timeline, avg_nnz, max_ctas, st);
  return static_cast<int>(cudaErrorInvalidValue);
}
