// lr_xchg.cu — BSP sparse-LR iterations whose Get / Add / Clock are ONE exchange per step:
// reduce at the owner, then the owner broadcasts.
//
// Reference semantics (server/consistency/bsp_model.cpp:14-43): adds of clock p are buffered and
// applied only when every worker has clocked; a Get of clock p+1 is answered after that.  The
// generic device path does it as pull -> compute -> push -> clock with three cross-GPU waits per
// iteration and the whole table crossing NVLink twice (lr.cu).  Here every rank keeps
//   replica   a full copy of the weights, kept current by the shard owners,
//   gacc[2]   its own gradient of the running batch (double buffered),
// in a symmetric buffer that every GPU maps (and, with NVLS, the switch maps as one multicast
// address).  Iteration p on every rank:
//   A  gate    folded[s] >= p for every shard (local spin)         = BSP Get gate
//      zero gacc[(p+1)&1]  (its last readers finished before folded reached p)
//   B  rows    SpMV + sigmoid + gradient of the batch -> gacc[p&1]        --- grid barrier
//   C  Clock   progress[worker] = p+1 on every rank (fence.sys first)
//   D  owner   wait min(progress) >= p+1; for every 16-byte chunk of MY key range:
//                 sum = multimem.ld_reduce(gacc[p&1])   (the switch adds the 8 GPUs' copies)
//                 w   = weights[chunk] + sum -> weights (the shard itself, master copy)
//                 multimem.st(replica[chunk], w)        (lands in all 8 replicas)
//              without NVLS: N peer loads and N peer stores per chunk instead.
//   E  fold    last CTA out: fence.sys, folded[my shards] = p+1 on every rank
// One data crossing of |table|/N bytes per direction and GPU with NVLS (the old path: 7/8 of the
// table each way), two flag waits, one grid barrier.  The flags are the generic tracker's
// (progress / folded), so generic Get / Add / Clock calls before and after a fused run see a
// consistent table.
//
// Requires: BSP, every rank of the table runs exactly one worker through this kernel with the
// same n_iters, one value per key.
#include "lr_common.cuh"

namespace mps {

template <typename T>
struct XchgPtrs {
  T* replica;        // local
  T* gacc;           // local, two buffers `stride` apart
  T* replica_mc;     // multicast or nullptr
  T* gacc_mc;
  int64_t stride;
};

// owner phase over flat elements [ob, oe) of the table, chunk-wise
template <typename T>
__device__ __forceinline__ void owner_reduce_bcast(const MpsTableView& v, const MpsClockView& c, const MpsXchg& x,
                                                   const XchgPtrs<T>& q, int buf, int64_t ob, int64_t oe,
                                                   int64_t tid, int64_t nthreads) {
  constexpr int N = Chunk<T>::N;
  const int R = c.num_ranks;
  const int64_t goff = static_cast<int64_t>(buf) * q.stride;
  const bool aligned = shards_chunk_aligned<T>(v) && (ob % N) == 0;
  if (aligned) {
    const int64_t nc = (oe - ob) / N;
    for (int64_t i = tid; i < nc; i += nthreads) {
      const int64_t f = ob + i * N;
      Chunk<T> sum;
      if (q.gacc_mc) {
        sum = mc_ld_reduce_chunk(q.gacc_mc + goff + f);
      } else {
        sum = ld_cg_chunk(q.gacc + goff + f);
        for (int k = 1; k < R; ++k) {
          int r = c.my_rank + k;
          if (r >= R) r -= R;
          sum.add(ld_sys_chunk(static_cast<const T*>(x.gacc[r]) + goff + f));
        }
      }
      T* wp = shard_elem<T>(v, f);
      Chunk<T> w;
      w.v = *reinterpret_cast<const decltype(w.v)*>(wp);
      w.add(sum);
      st_chunk(wp, w);
      if (q.replica_mc) {
        mc_st_chunk(q.replica_mc + f, w);
      } else {
        st_chunk(q.replica + f, w);
        for (int k = 1; k < R; ++k) {
          int r = c.my_rank + k;
          if (r >= R) r -= R;
          st_sys_chunk(static_cast<T*>(x.replica[r]) + f, w);
        }
      }
    }
    ob += nc * N;  // tail below
  }
  for (int64_t f = ob + tid; f < oe; f += nthreads) {
    T sum;
    if (q.gacc_mc) {
      sum = mc_ld_reduce(q.gacc_mc + goff + f);
    } else {
      sum = ld_cg(q.gacc + goff + f);
      for (int k = 1; k < R; ++k) {
        int r = c.my_rank + k;
        if (r >= R) r -= R;
        sum += ld_sys(static_cast<const T*>(x.gacc[r]) + goff + f);
      }
    }
    T* wp = shard_elem<T>(v, f);
    const T w = *wp + sum;
    *wp = w;
    if (q.replica_mc) {
      mc_st(q.replica_mc + f, w);
    } else {
      q.replica[f] = w;
      for (int k = 1; k < R; ++k) {
        int r = c.my_rank + k;
        if (r >= R) r -= R;
        st_sys(static_cast<T*>(x.replica[r]) + f, w);
      }
    }
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

template <int L, typename T>
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
  int* pad = c.pad[c.my_rank];
  GridBarrier grid{pad + MPS_PAD_LOCALCTR + 2, 0};  // zeroed by the host before the launch
  const bool tl = timeline != nullptr && tid == 0;
  XchgPtrs<T> q;
  q.replica = static_cast<T*>(x.replica[c.my_rank]);
  q.gacc = static_cast<T*>(x.gacc[c.my_rank]);
  q.replica_mc = static_cast<T*>(x.replica_mc);
  q.gacc_mc = static_cast<T*>(x.gacc_mc);
  q.stride = x.gacc_stride;
  const int64_t ob = c.own_count > 0 ? flat_begin1(v, c.own_first) : 0;
  const int64_t oe = c.own_count > 0 ? flat_begin1(v, c.own_first + c.own_count) : 0;

  if (!x.replica_valid) {
    // First launch on this table (or the weights were changed behind the replica's back): wait
    // for the Get gate of start_clock, copy every shard once, start from clean accumulators.
    if (threadIdx.x == 0) spin_until(c, [&] { return gate_open(c, start_clock); });
    __syncthreads();
    replica_fill<T>(v, c.own_first, q.replica, tid, nthreads);
    zero_range(q.gacc + static_cast<int64_t>(start_clock & 1) * q.stride, static_cast<int64_t>(n_keys), tid,
               nthreads);
    grid.sync();
  }

  for (int it = 0; it < n_iters; ++it) {
    const int p = start_clock + it;
    const int buf = p & 1;
    if (tl) timeline[8 * it + 0] = globaltimer_ns();
    // A. gate: every shard folded up to p (owners have written the weights of clock p into my
    //    replica and are done reading gacc[buf ^ 1])
    //    (a timed-out wait raises the pad error flag and every later wait falls through, so a lost
    //    peer ends the launch with an error on the host instead of a hung GPU)
    if (threadIdx.x == 0) spin_until(c, [&] { return gate_open(c, p); });
    __syncthreads();
    if (tl) timeline[8 * it + 1] = globaltimer_ns();
    zero_range(q.gacc + static_cast<int64_t>(buf ^ 1) * q.stride, static_cast<int64_t>(n_keys), tid, nthreads);
    // B. rows
    float loss = 0.f, correct = 0.f;
    rows_pass<L, true, T>(csr, batch_starts[it], batch, key0, n_keys, q.replica,
                          q.gacc + static_cast<int64_t>(buf) * q.stride, alpha, multi, tid / L, nthreads / L, loss,
                          correct);
    block_accumulate(loss, correct, loss_out + 2 * it);
    grid.sync();
    if (tl) timeline[8 * it + 2] = globaltimer_ns();
    // C. Clock: my gradient of clock p is complete.  The grid barrier ordered every CTA's reds
    //    before this thread; the fence makes them visible system-wide before the flag.
    if (tid == 0) {
      if (multi) fence_sys(); else __threadfence();
      for (int i = 0; i < c.num_ranks; ++i) {
        const int r = (c.my_rank + i) % c.num_ranks;
        st_relaxed_sys(c.pad[r] + MPS_PAD_PROGRESS + worker, p + 1);
      }
    }
    // D. owner phase
    if (oe > ob) {
      if (threadIdx.x == 0) spin_until(c, [&] { return pad_min_clock(c) >= p + 1; });
      __syncthreads();
      if (tl) timeline[8 * it + 3] = globaltimer_ns();
      owner_reduce_bcast<T>(v, c, x, q, buf, ob, oe, tid, nthreads);
      if (tl) timeline[8 * it + 4] = globaltimer_ns();
      // E. last CTA out publishes folded[s] = p+1.  Every CTA fences its own stores at system
      //    scope before it arrives; the last one fences again before the flag stores.
      __syncthreads();
      if (threadIdx.x == 0) {
        if (multi) fence_sys(); else __threadfence();
        const int done = atomicAdd(pad + MPS_PAD_LOCALCTR + 1, 1) + 1;
        if (done == static_cast<int>(gridDim.x)) {
          pad[MPS_PAD_LOCALCTR + 1] = 0;
          if (multi) fence_sys(); else __threadfence();
          for (int i = 0; i < c.num_ranks; ++i) {
            const int r = (c.my_rank + i) % c.num_ranks;
            for (int s = c.own_first; s < c.own_first + c.own_count; ++s)
              st_relaxed_sys(c.pad[r] + MPS_PAD_FOLDED + s, p + 1);
          }
        }
      }
    }
    if (tl) timeline[8 * it + 5] = globaltimer_ns();
  }
}

template <typename T>
static void* xchg_kernel_for(int L) {
  return L == 8 ? reinterpret_cast<void*>(lr_train_xchg_kernel<8, T>)
                : (L == 16 ? reinterpret_cast<void*>(lr_train_xchg_kernel<16, T>)
                           : reinterpret_cast<void*>(lr_train_xchg_kernel<32, T>));
}

template <typename T>
static int launch_xchg(const MpsTableView* view, const MpsClockView* clk, int worker, int start_clock,
                       const MpsCsr* csr, const int64_t* batch_starts, int n_iters, int64_t batch, double alpha_f,
                       const MpsXchg* xchg, float* loss_out, unsigned long long* timeline, int avg_nnz,
                       int max_ctas, cudaStream_t st) {
  int dev = 0, sms = 148, per_sm = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int L = lanes_for(avg_nnz > 0 ? avg_nnz : 40);
  void* fn = xchg_kernel_for<T>(L);
  MPS_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, 512, 0));
  if (per_sm < 1) return static_cast<int>(cudaErrorLaunchOutOfResources);
  int grid = sms * per_sm;
  if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
  MpsTableView v = *view;
  MpsClockView c = *clk;
  MpsCsr m = *csr;
  MpsXchg x = *xchg;
  T alpha = static_cast<T>(alpha_f);
  void* args[] = {&v, &c, &worker, &start_clock, &m, &batch_starts, &n_iters, &batch, &alpha, &x, &loss_out,
                  &timeline};
  MPS_CHECK(cudaMemsetAsync(c.pad[c.my_rank] + MPS_PAD_LOCALCTR + 2, 0, sizeof(int), st));
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
