// lr_bucket.cu — sparse-regime LR step (batch touches a small part of a huge table, e.g. the 100M-key
// model) as a request / reply / delta MESSAGE flow between GPUs instead of one small NVLink packet per
// non-zero.
//
// The reference's iteration is exactly that message flow (worker/kv_client_table.hpp:168-195: Get = keys
// sliced per server -> kGet -> values back; Add = keys + deltas per server -> kAdd, applied by the server
// thread).  The first device version (lr.cu, lr_step_sparse_kernel) replaced it by one peer `ld` and one
// peer `red` per non-zero; with 8 GPUs that is 4.6 M scattered 4-byte packets per GPU and step, and it
// runs at the small-packet rate of the fabric (~5 % of the link).  Here every GPU plays both roles inside
// one cooperative kernel per step:
//   1  bucket   the batch's feature ids are counted and scattered into per-owner buckets (local), every
//               non-zero remembers its slot
//   2  request  the buckets go to their owners' inboxes as contiguous, fully coalesced peer stores
//   3  serve    every owner gathers the requested weights from its own HBM and writes them, again as one
//               contiguous stream, into the requester's reply buffer
//   4  compute  forward SpMV + sigmoid + gradient against the LOCAL reply buffer; deltas written per slot
//   5  delta    the delta buckets go to the owners (same positions as the keys of step 2)
//   6  apply    every owner adds what it received into its shard with local atomics
// NVLink carries 3 x 4 bytes per non-zero in long bursts; all random access is local.  Three grid
// barriers and three cross-GPU waits per step (per-CTA counters behind one system fence each, as in
// lr_xchg.cu).  Adds of a step are applied by an owner only after it has served every request of that
// step, so the flow is BSP-tight; it also satisfies SSP / ASP, whose bounds are upper bounds.
//
// Requires: every rank runs exactly one worker through this kernel, step by step, with the same grid
// size; non-wrapping batches; fp32 or fp64 tables with one value per key.
#include "lr_common.cuh"

namespace mps {

constexpr int kBkThreads = 512;

template <typename T>
struct BucketPtrs {
  uint32_t* inbox_keys;  // [R][cap]   keys requested by rank s (local copy of my inbox)
  T* inbox_delta;        // [R][cap]
  int64_t* inbox_meta;   // [R][2]     {count, offset of this bucket in the requester's buffers}
  T* reply;              // [cap_total] values for MY requests, bucket order
};

template <typename T>
__device__ __forceinline__ BucketPtrs<T> bucket_ptrs(const MpsBucket& b, int rank) {
  char* base = static_cast<char*>(b.base[rank]);
  BucketPtrs<T> p;
  p.inbox_keys = reinterpret_cast<uint32_t*>(base + b.off_inbox_keys);
  p.inbox_delta = reinterpret_cast<T*>(base + b.off_inbox_delta);
  p.inbox_meta = reinterpret_cast<int64_t*>(base + b.off_inbox_meta);
  p.reply = reinterpret_cast<T*>(base + b.off_reply);
  return p;
}

__device__ __forceinline__ void st_sys_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_sys_s64(int64_t* p, int64_t v) {
  asm volatile("st.relaxed.sys.global.s64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ int64_t ld_sys_s64(const int64_t* p) {
  int64_t v;
  asm volatile("ld.relaxed.sys.global.s64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// rank that owns `key` (shards are laid out rank-major, the same number on every rank), or -1
__device__ __forceinline__ int owner_rank(const MpsTableView& v, uint32_t key, int shards_per_rank) {
  const int s = owner_of(v, key);
  return s < 0 ? -1 : s / shards_per_rank;
}

template <int L, typename T>
__global__ void __launch_bounds__(kBkThreads, 2)
lr_step_bucket_kernel(const __grid_constant__ MpsTableView v, const __grid_constant__ MpsClockView c, int gated,
                      int progress, const __grid_constant__ MpsCsr csr, int64_t row_begin, int64_t batch, T alpha,
                      const __grid_constant__ MpsBucket b, float* __restrict__ loss_out) {
  const int64_t tid = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int64_t nthreads = static_cast<int64_t>(gridDim.x) * blockDim.x;
  const int G = static_cast<int>(gridDim.x);
  const int R = c.num_ranks, me = c.my_rank;
  const int spr = v.num_shards / R;
  const bool multi = R > 1;
  int* pad = c.pad[me];
  GridBarrier grid{pad + MPS_PAD_LOCALCTR + 2, 0};  // zeroed by the host before the launch
  const int set = MPS_PAD_BCNT + (b.seq & 1) * 4 * MPS_MAX_RANKS;
  const int kReq = set, kRep = set + MPS_MAX_RANKS, kDlt = set + 2 * MPS_MAX_RANKS, kApl = set + 3 * MPS_MAX_RANKS;
  const int64_t nnz0 = csr.row_ptr[row_begin];
  const int64_t M = csr.row_ptr[row_begin + batch] - nnz0;
  const uint32_t* col = csr.col + nnz0;
  const float* val = csr.val + nnz0;
  uint32_t* send_keys = b.send_keys;
  uint32_t* slot = b.slot;
  T* delta_send = static_cast<T*>(b.delta_send);
  unsigned long long* ghist = b.hist;  // [R] totals, zeroed by the host before the launch
  const BucketPtrs<T> mine = bucket_ptrs<T>(b, me);
  __shared__ unsigned int s_cnt[MPS_MAX_RANKS];
  __shared__ unsigned long long s_base[MPS_MAX_RANKS];
  __shared__ long long s_off[MPS_MAX_RANKS + 1];

  if (M > b.cap) {  // the host sized the buffers for less: refuse (uniformly, before any barrier)
    if (tid == 0) st_relaxed_sys(pad + MPS_PAD_ERROR, 1);
    return;
  }
  const bool tl = b.timeline != nullptr && tid == 0;
  if (tl) b.timeline[0] = globaltimer_ns();
  // The counter set of the NEXT launch is cleared now: its signals of two launches ago were all consumed
  // before that launch ended, and no peer can start the next launch before this one has finished here.
  if (blockIdx.x == 0 && threadIdx.x < 4 * MPS_MAX_RANKS)
    pad[MPS_PAD_BCNT + ((b.seq + 1) & 1) * 4 * MPS_MAX_RANKS + threadIdx.x] = 0;
  if (gated) cta_gate(c, progress);
  if (tl) b.timeline[1] = globaltimer_ns();

  // ---- 1. bucket: count, then scatter.  Every CTA owns a contiguous range of the batch's non-zeros.
  const int64_t chunk = (M + G - 1) / G;
  const int64_t j0 = static_cast<int64_t>(blockIdx.x) * chunk;
  const int64_t j1 = j0 + chunk < M ? j0 + chunk : M;
  if (threadIdx.x < MPS_MAX_RANKS) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  // One shared-memory atomic per distinct owner and warp (match_any), not one per non-zero; the owner
  // found in the counting pass is parked in `slot` for the scatter pass.
  // Four independent loads per thread are issued before the first is used (the loop is latency-bound
  // otherwise: match_any is a convergence point the compiler will not move loads across).
  const unsigned lane = threadIdx.x & 31;
  constexpr int U = 4;
  for (int64_t j = j0 + threadIdx.x; j < j1; j += U * blockDim.x) {
    uint32_t key[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t jj = j + static_cast<int64_t>(u) * blockDim.x;
      key[u] = jj < j1 ? ld_stream(col + jj) : 0u;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t jj = j + static_cast<int64_t>(u) * blockDim.x;
      if (jj < j1) {
        const int o = owner_rank(v, key[u], spr);
        slot[jj] = static_cast<uint32_t>(o);
        const unsigned grp = __match_any_sync(__activemask(), o);
        if (o >= 0 && lane == static_cast<unsigned>(__ffs(grp) - 1)) atomicAdd(&s_cnt[o], __popc(grp));
      }
    }
  }
  __syncthreads();
  if (threadIdx.x < R) {
    s_base[threadIdx.x] = atomicAdd(ghist + threadIdx.x, static_cast<unsigned long long>(s_cnt[threadIdx.x]));
    s_cnt[threadIdx.x] = 0;
  }
  grid.sync();
  if (threadIdx.x == 0) {
    long long acc = 0;
    for (int r = 0; r < R; ++r) {
      s_off[r] = acc;
      acc += static_cast<long long>(*reinterpret_cast<volatile unsigned long long*>(ghist + r));
    }
    s_off[R] = acc;
  }
  __syncthreads();
  for (int64_t j = j0 + threadIdx.x; j < j1; j += U * blockDim.x) {
    uint32_t key[U];
    int own[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t jj = j + static_cast<int64_t>(u) * blockDim.x;
      key[u] = jj < j1 ? ld_stream(col + jj) : 0u;
      own[u] = jj < j1 ? static_cast<int>(slot[jj]) : -1;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t jj = j + static_cast<int64_t>(u) * blockDim.x;
      if (jj < j1) {
        const int o = own[u];
        const unsigned grp = __match_any_sync(__activemask(), o);
        uint32_t sl = 0xffffffffu;
        if (o >= 0) {
          const int leader = __ffs(grp) - 1;
          unsigned base = 0;
          if (lane == static_cast<unsigned>(leader)) base = atomicAdd(&s_cnt[o], __popc(grp));
          base = __shfl_sync(grp, base, leader);
          const long long pos =
              s_off[o] + static_cast<long long>(s_base[o]) + base + __popc(grp & ((1u << lane) - 1u));
          sl = static_cast<uint32_t>(pos);
          send_keys[pos] = key[u];
        }
        slot[jj] = sl;
      }
    }
  }
  grid.sync();
  // every CTA has its offsets: leave the totals at zero for the next launch
  if (blockIdx.x == 0 && threadIdx.x < MPS_MAX_RANKS) ghist[threadIdx.x] = 0;
  if (tl) b.timeline[2] = globaltimer_ns();

  // ---- 2. request: every bucket to its owner's inbox, contiguous peer stores
  for (int k = 0; k < R; ++k) {
    const int o = (me + 1 + k) % R;
    const long long beg = s_off[o], cnt = s_off[o + 1] - beg;
    uint32_t* dst = bucket_ptrs<T>(b, o).inbox_keys + static_cast<int64_t>(me) * b.cap;
    // 16-byte peer stores (the inbox is aligned, the bucket inside send_keys need not be: 4 scalar loads)
    const int64_t nq = cnt >> 2;
    for (int64_t q = tid; q < nq; q += nthreads) {
      const uint32_t* sp = send_keys + beg + 4 * q;
      asm volatile("st.relaxed.sys.global.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(dst + 4 * q), "r"(sp[0]), "r"(sp[1]),
                   "r"(sp[2]), "r"(sp[3])
                   : "memory");
    }
    for (int64_t i = (nq << 2) + tid; i < cnt; i += nthreads) st_sys_u32(dst + i, send_keys[beg + i]);
    if (tid == 0) {
      int64_t* meta = bucket_ptrs<T>(b, o).inbox_meta + 2 * me;
      st_sys_s64(meta, cnt);
      st_sys_s64(meta + 1, beg);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    cta_signal(c, kReq, multi);
    spin_until(c, [&] { return all_reached(pad + kReq, R, G); });
  }
  __syncthreads();
  if (tl) b.timeline[3] = globaltimer_ns();

  // ---- 3. serve: gather from my own shards, reply as one contiguous stream per requester
  for (int k = 0; k < R; ++k) {
    const int s = (me + 1 + k) % R;
    const int64_t cnt = ld_sys_s64(mine.inbox_meta + 2 * s), roff = ld_sys_s64(mine.inbox_meta + 2 * s + 1);
    const uint32_t* keys = mine.inbox_keys + static_cast<int64_t>(s) * b.cap;
    T* dst = bucket_ptrs<T>(b, s).reply + roff;
    for (int64_t i = tid; i < cnt; i += U * nthreads) {
      uint32_t key[U];
      T w[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t ii = i + u * nthreads;
        key[u] = ii < cnt ? ld_sys_u32(keys + ii) : v.begin[c.own_first];
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int sh = owner_of(v, key[u]);
        w[u] = ld_cg(static_cast<const T*>(v.weights[sh]) + (key[u] - v.begin[sh]));
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t ii = i + u * nthreads;
        if (ii < cnt) st_sys(dst + ii, w[u]);
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    cta_signal(c, kRep, multi);
    spin_until(c, [&] { return all_reached(pad + kRep, R, G); });
  }
  __syncthreads();
  if (tl) b.timeline[4] = globaltimer_ns();

  // ---- 4. compute: rows against the local reply buffer; one delta per slot (no atomics)
  {
    const int lane = threadIdx.x & 31;
    const int gl = lane & (L - 1);
    const unsigned mask = (L == 32) ? 0xffffffffu : (((1u << L) - 1u) << (lane & ~(L - 1)));
    const int64_t group = tid / L, ngroups = nthreads / L;
    float loss = 0.f, correct = 0.f;
    for (int64_t r = group; r < batch; r += ngroups) {
      const int64_t ridx = row_begin + r;
      const int64_t beg = csr.row_ptr[ridx] - nnz0, end = csr.row_ptr[ridx + 1] - nnz0;
      // up to E non-zeros per lane live in registers: slots, then replies, then values are each loaded
      // as a batch; the same registers feed the delta pass
      constexpr int E = kRowRegs;
      uint32_t sl[E];
      T x[E], w[E];
#pragma unroll
      for (int e = 0; e < E; ++e) {
        const int64_t j = beg + gl + e * L;
        sl[e] = j < end ? slot[j] : 0xffffffffu;
        x[e] = j < end ? static_cast<T>(ld_stream(val + j)) : T(0);
      }
#pragma unroll
      for (int e = 0; e < E; ++e) w[e] = sl[e] != 0xffffffffu ? ld_cg(mine.reply + sl[e]) : T(0);
      T acc = T(0);
#pragma unroll
      for (int e = 0; e < E; ++e) acc = fma(w[e], x[e], acc);
      for (int64_t j = beg + gl + E * L; j < end; j += L) {  // rows longer than L*E non-zeros
        const uint32_t s2 = slot[j];
        if (s2 != 0xffffffffu) acc = fma(ld_cg(mine.reply + s2), static_cast<T>(ld_stream(val + j)), acc);
      }
      const T z = group_sum<L>(acc, mask);
      const T y = static_cast<T>(csr.label[ridx]);
      const T pred = sigmoid_t(z);
      const T g = alpha * (y - pred);
      if (gl == 0) {
        loss += static_cast<float>(logloss_t(z, y));
        correct += ((pred > T(0.5)) == (y > T(0.5))) ? 1.f : 0.f;
      }
#pragma unroll
      for (int e = 0; e < E; ++e)
        if (sl[e] != 0xffffffffu) delta_send[sl[e]] = g * x[e];
      for (int64_t j = beg + gl + E * L; j < end; j += L) {
        const uint32_t s2 = slot[j];
        if (s2 != 0xffffffffu) delta_send[s2] = g * static_cast<T>(ld_stream(val + j));
      }
    }
    block_accumulate(loss, correct, loss_out);
  }
  grid.sync();
  if (tl) b.timeline[5] = globaltimer_ns();

  // ---- 5. delta: the delta buckets to the owners, same positions as the keys
  for (int k = 0; k < R; ++k) {
    const int o = (me + 1 + k) % R;
    const long long beg = s_off[o], cnt = s_off[o + 1] - beg;
    T* dst = bucket_ptrs<T>(b, o).inbox_delta + static_cast<int64_t>(me) * b.cap;
    constexpr int N = Chunk<T>::N;
    const int64_t nq = cnt / N;
    for (int64_t q = tid; q < nq; q += nthreads) {
      const T* sp = delta_send + beg + N * q;
      Chunk<T> ch;
      if constexpr (N == 4) ch.v = make_float4(sp[0], sp[1], sp[2], sp[3]);
      else ch.v = make_double2(sp[0], sp[1]);
      st_sys_chunk(dst + N * q, ch);
    }
    for (int64_t i = nq * N + tid; i < cnt; i += nthreads) st_sys(dst + i, delta_send[beg + i]);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    cta_signal(c, kDlt, multi);
    spin_until(c, [&] { return all_reached(pad + kDlt, R, G); });
  }
  __syncthreads();
  if (tl) b.timeline[6] = globaltimer_ns();

  // ---- 6. apply: server-side += (server/vector_storage.hpp:28-38) with local atomics
  for (int k = 0; k < R; ++k) {
    const int s = (me + 1 + k) % R;
    const int64_t cnt = ld_sys_s64(mine.inbox_meta + 2 * s);
    const uint32_t* keys = mine.inbox_keys + static_cast<int64_t>(s) * b.cap;
    const T* dl = mine.inbox_delta + static_cast<int64_t>(s) * b.cap;
    for (int64_t i = tid; i < cnt; i += U * nthreads) {
      uint32_t key[U];
      T d[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t ii = i + u * nthreads;
        d[u] = ii < cnt ? ld_cg(dl + ii) : T(0);
        key[u] = ii < cnt ? ld_sys_u32(keys + ii) : 0u;
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (d[u] == T(0)) continue;
        const int sh = owner_of(v, key[u]);
        red_gpu(static_cast<T*>(v.weights[sh]) + (key[u] - v.begin[sh]), d[u]);
      }
    }
  }
  if (tl) b.timeline[7] = globaltimer_ns();

  // ---- 7. Clock: once every owner has applied every delta of this step, the step is visible everywhere.
  //      Each rank records that in its OWN pad (every worker clocked, every shard folded): the generic
  //      tracker — a Get gate, a checkpoint — sees the run without a separate clock launch.
  __syncthreads();
  if (threadIdx.x == 0) {
    cta_signal(c, kApl, multi);
    spin_until(c, [&] { return all_reached(pad + kApl, R, G); });
  }
  if (tid == 0) {
    for (int w = 0; w < c.num_workers; ++w) st_relaxed_sys(pad + MPS_PAD_PROGRESS + w, progress + 1);
    for (int sh = 0; sh < c.num_shards; ++sh) st_relaxed_sys(pad + MPS_PAD_FOLDED + sh, progress + 1);
    pad[MPS_PAD_LOCALCTR + 2] = 0;  // grid-barrier counter: every local CTA is past its last barrier
  }
}

template <typename T>
static int launch_bucket(const MpsTableView* view, const MpsClockView* clk, int gated, int progress,
                         const MpsCsr* csr, int64_t row_begin, int64_t batch, double alpha_f, const MpsBucket* bk,
                         float* loss_out, int avg_nnz, cudaStream_t st) {
  int dev = 0, sms = 148, per_sm = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int L = lanes_for(avg_nnz > 0 ? avg_nnz : 40);
  void* fn = L == 8 ? reinterpret_cast<void*>(lr_step_bucket_kernel<8, T>)
                    : (L == 16 ? reinterpret_cast<void*>(lr_step_bucket_kernel<16, T>)
                               : reinterpret_cast<void*>(lr_step_bucket_kernel<32, T>));
  MPS_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, kBkThreads, 0));
  if (per_sm < 1) return static_cast<int>(cudaErrorLaunchOutOfResources);
  const int grid = sms * per_sm;
  MpsTableView v = *view;
  MpsClockView c = *clk;
  MpsCsr m = *csr;
  MpsBucket b = *bk;
  T alpha = static_cast<T>(alpha_f);
  void* args[] = {&v, &c, &gated, &progress, &m, &row_begin, &batch, &alpha, &b, &loss_out};
  // no memsets: the kernel leaves the grid-barrier counter and the bucket totals at zero and clears the
  // next launch's counter set itself (reset_pad zeroes everything when a task starts)
  MPS_CHECK(cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(kBkThreads), args, 0, st));
  return 0;
}

}  // namespace mps

extern "C" int mps_lr_step_bucketed(const MpsTableView* view, const MpsClockView* clk, int gated, int progress,
                                    const MpsCsr* csr, int64_t row_begin, int64_t batch, double alpha,
                                    const MpsBucket* bucket, float* loss_out, int avg_nnz, void* stream) {
  if (batch <= 0) return 0;
  if (view->vals_per_key != 1 || csr->format != MPS_CSR_PLAIN || row_begin < 0 || row_begin + batch > csr->n_rows ||
      view->num_shards % clk->num_ranks != 0)
    return static_cast<int>(cudaErrorInvalidValue);
  auto st = static_cast<cudaStream_t>(stream);
  if (view->dtype == MPS_F64)
    return mps::launch_bucket<double>(view, clk, gated, progress, csr, row_begin, batch, alpha, bucket, loss_out,
                                      avg_nnz, st);
  if (view->dtype == MPS_F32)
    return mps::launch_bucket<float>(view, clk, gated, progress, csr, row_begin, batch, alpha, bucket, loss_out,
                                     avg_nnz, st);
  return static_cast<int>(cudaErrorInvalidValue);
}
