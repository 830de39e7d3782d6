// clock.cu — KVClientTable::Clock, the progress tracker, the BSP clock edge and the
// process barrier as device-side flags in peer memory (no server thread, no message).
//
// Reference: kClock broadcast to every server thread (worker/kv_client_table.hpp:123-134),
// ProgressTracker::AdvanceAndGetChangedMinClock (server/util/progress_tracker.cpp:46-72),
// BSPModel::Clock applying buffered adds then releasing buffered gets
// (server/consistency/bsp_model.cpp:14-32), Mailbox::Barrier (comm/mailbox.cpp:362-396).
//
// Protocol.  Every rank keeps a replica of the progress table in its signal pad; a worker's
// Clock stores its new clock into slot `worker` of EVERY replica with release.sys, so gates
// spin on local memory only.  Under BSP the owner of a shard waits for the edge (all
// workers >= t), folds staging into weights, clears staging and publishes folded[s] = t
// to every replica; a Get at clock t waits for folded[s] >= t of every shard.
#include "common.cuh"

namespace mps {

// With a single rank every reader of the pad and the shards is on this device: gpu-scope fences
// are enough and cost a fraction of MEMBAR.SYS.
__device__ __forceinline__ void fence_for(const MpsClockView& c) {
  if (c.num_ranks > 1) fence_sys(); else __threadfence();
}

__device__ __forceinline__ void clock_body(const MpsTableView& v, const MpsClockView& c, int worker, int new_clock,
                                           int fold, int bx, int nb) {
  // (1) signal.  Kernel boundary ordered this device's earlier pushes; the fence makes them
  //     visible system-wide before any replica can observe the new clock.
  if (worker >= 0 && bx == 0 && threadIdx.x == 0) {
    fence_for(c);
    for (int i = 0; i < c.num_ranks; ++i) {
      const int r = (c.my_rank + i) % c.num_ranks;
      st_relaxed_sys(c.pad[r] + MPS_PAD_PROGRESS + worker, new_clock);
    }
  }
  if (!fold || c.model != MPS_MODEL_BSP || c.own_count <= 0) return;

  // (2) clock edge: wait until every worker has clocked past new_clock - 1.
  __shared__ int ok;
  if (threadIdx.x == 0) ok = spin_until(c, [&] { return pad_min_clock(c) >= new_clock; }) ? 1 : 0;
  __syncthreads();
  if (!ok) return;

  // (3) fold my shards: weights += staging; staging = 0 (apply add_buffer_, bsp_model.cpp:20-23)
  const int64_t tid = static_cast<int64_t>(bx) * blockDim.x + threadIdx.x;
  const int64_t stride = static_cast<int64_t>(nb) * blockDim.x;
  for (int s = c.own_first; s < c.own_first + c.own_count; ++s) {
    if (!v.staging[s]) continue;
    const int64_t len = static_cast<int64_t>(v.begin[s + 1] - v.begin[s]) * v.vals_per_key;
    if (v.dtype == MPS_F64) {
      double* w = static_cast<double*>(v.weights[s]);
      double* st = static_cast<double*>(v.staging[s]);
      for (int64_t i = tid; i < len; i += stride) {
        const double d = st[i];
        if (d != 0.0) { w[i] += d; st[i] = 0.0; }
      }
    } else {
      float* w = static_cast<float*>(v.weights[s]);
      float* st = static_cast<float*>(v.staging[s]);
      const int64_t len4 = len >> 2;
      for (int64_t i = tid; i < len4; i += stride) {
        const float4 d = ld_cg_v4(st + 4 * i);
        if (d.x != 0.f || d.y != 0.f || d.z != 0.f || d.w != 0.f) {
          float4 x = *reinterpret_cast<float4*>(w + 4 * i);
          x.x += d.x; x.y += d.y; x.z += d.z; x.w += d.w;
          *reinterpret_cast<float4*>(w + 4 * i) = x;
          *reinterpret_cast<float4*>(st + 4 * i) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
      for (int64_t i = (len4 << 2) + tid; i < len; i += stride) {
        const float d = ld_cg(st + i);
        if (d != 0.f) { w[i] += d; st[i] = 0.f; }
      }
    }
  }

  // (4) last CTA out publishes folded[s] = new_clock on every replica.
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    int* ctr = c.pad[c.my_rank] + MPS_PAD_LOCALCTR;
    const int done = atomicAdd(ctr, 1) + 1;
    if (done == nb) {
      *ctr = 0;
      fence_for(c);
      for (int i = 0; i < c.num_ranks; ++i) {
        const int r = (c.my_rank + i) % c.num_ranks;
        for (int s = c.own_first; s < c.own_first + c.own_count; ++s)
          st_relaxed_sys(c.pad[r] + MPS_PAD_FOLDED + s, new_clock);
      }
    }
  }
}

__global__ void __launch_bounds__(256)
clock_kernel(const __grid_constant__ MpsTableView v, const __grid_constant__ MpsClockView c,
             int worker, int new_clock, int fold) {
  clock_body(v, c, worker, new_clock, fold, blockIdx.x, gridDim.x);
}

// Clock of two tables of one worker in a single launch (k-means: centres + counts): blockIdx.y picks
// the table, nb0 / nb1 CTAs take part in each fold.
__global__ void __launch_bounds__(256)
clock2_kernel(const __grid_constant__ MpsTableView v0, const __grid_constant__ MpsClockView c0, int nb0,
              const __grid_constant__ MpsTableView v1, const __grid_constant__ MpsClockView c1, int nb1,
              int worker, int new_clock, int fold) {
  if (blockIdx.y == 0) {
    if (static_cast<int>(blockIdx.x) < nb0) clock_body(v0, c0, worker, new_clock, fold, blockIdx.x, nb0);
  } else {
    if (static_cast<int>(blockIdx.x) < nb1) clock_body(v1, c1, worker, new_clock, fold, blockIdx.x, nb1);
  }
}

// Resident server of a rank that owns BSP shards but runs no worker of the task (the reference's server
// thread applies the buffered adds at every clock edge, bsp_model.cpp:14-32): loops
//   wait until every worker clocked past `next`  ->  fold my shards  ->  publish folded = next
// until the host raises *done (after the workers' ranks reported the end of the task) and no further
// edge is pending.  No host round trip gates an edge.  Idle waiting is not an error: only the host ends it.
__global__ void __launch_bounds__(256)
fold_server_kernel(const __grid_constant__ MpsTableView v, const __grid_constant__ MpsClockView c, int start_clock,
                   const volatile int* __restrict__ done) {
  __shared__ int s_go;
  int next = start_clock + 1;
  while (true) {
    if (threadIdx.x == 0) {
      unsigned ns = 64;
      while (true) {
        if (pad_min_clock(c) >= next) { s_go = 1; break; }
        if (*done != 0 && pad_min_clock(c) < next) { s_go = 0; break; }
        if (pad_error(c)) { s_go = 0; break; }
        __nanosleep(ns);
        if (ns < 2048) ns <<= 1;
      }
    }
    __syncthreads();
    const int go = s_go;
    __syncthreads();
    if (!go) return;
    clock_body(v, c, -1, next, 1, blockIdx.x, gridDim.x);
    ++next;
  }
}

__global__ void barrier_kernel(const __grid_constant__ MpsClockView c, int epoch) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  fence_sys();
  for (int i = 0; i < c.num_ranks; ++i) {
    const int r = (c.my_rank + i) % c.num_ranks;
    st_relaxed_sys(c.pad[r] + MPS_PAD_BARRIER + c.my_rank, epoch);
  }
  const int* bar = c.pad[c.my_rank] + MPS_PAD_BARRIER;
  spin_until(c, [&] {
    for (int r = 0; r < c.num_ranks; ++r)
      if (ld_acquire_sys(bar + r) < epoch) return false;
    return true;
  });
}

// Device-side straggler injection (the reference sleeps the worker thread, lr_example.cpp:347-353): hold
// the worker's stream for `ns` nanoseconds so that its clock really lags on the device timeline.
__global__ void straggle_kernel(long long ns) {
  unsigned long long t0, t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  do {
    __nanosleep(1000);
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  } while (static_cast<long long>(t - t0) < ns);
}

__global__ void reset_pad_kernel(const __grid_constant__ MpsClockView c, int start_clock) {
  int* pad = c.pad[c.my_rank];
  for (int i = threadIdx.x; i < MPS_MAX_WORKERS; i += blockDim.x) {
    pad[MPS_PAD_PROGRESS + i] = start_clock;
    pad[MPS_PAD_PULLED + i] = start_clock;
  }
  for (int i = threadIdx.x; i < MPS_MAX_SHARDS; i += blockDim.x) pad[MPS_PAD_FOLDED + i] = start_clock;
  if (threadIdx.x == 0) {
    pad[MPS_PAD_ERROR] = 0;
    pad[MPS_PAD_MINCLOCK] = start_clock;
    for (int i = 0; i < 4; ++i) pad[MPS_PAD_LOCALCTR + i] = 0;
  }
  for (int i = threadIdx.x; i < 4 * MPS_MAX_RANKS; i += blockDim.x) pad[MPS_PAD_XCNT + i] = 0;
  for (int i = threadIdx.x; i < 8 * MPS_MAX_RANKS; i += blockDim.x) pad[MPS_PAD_BCNT + i] = 0;
}

}  // namespace mps

extern "C" {

static int clock_grid(const MpsTableView* view, const MpsClockView* clk, int fold) {
  if (!(fold && clk->model == MPS_MODEL_BSP && clk->own_count > 0)) return 1;
  int64_t len = 0;
  for (int s = clk->own_first; s < clk->own_first + clk->own_count; ++s)
    len += static_cast<int64_t>(view->begin[s + 1] - view->begin[s]) * view->vals_per_key;
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  }
  const int64_t want = (len / 4 + 256 * 4 - 1) / (256 * 4);
  return static_cast<int>(want < 1 ? 1 : (want > sms * 4 ? sms * 4 : want));
}

int mps_clock(const MpsTableView* view, const MpsClockView* clk, int worker, int new_clock, int fold,
              void* stream) {
  auto st = static_cast<cudaStream_t>(stream);
  const int grid = clock_grid(view, clk, fold);
  mps::clock_kernel<<<grid, 256, 0, st>>>(*view, *clk, worker, new_clock, fold);
  MPS_LAUNCH_CHECK();
  return 0;
}

int mps_clock2(const MpsTableView* view0, const MpsClockView* clk0, const MpsTableView* view1,
               const MpsClockView* clk1, int worker, int new_clock, int fold, void* stream) {
  auto st = static_cast<cudaStream_t>(stream);
  const int g0 = clock_grid(view0, clk0, fold), g1 = clock_grid(view1, clk1, fold);
  mps::clock2_kernel<<<dim3(g0 > g1 ? g0 : g1, 2), 256, 0, st>>>(*view0, *clk0, g0, *view1, *clk1, g1, worker,
                                                                  new_clock, fold);
  MPS_LAUNCH_CHECK();
  return 0;
}

int mps_fold_server(const MpsTableView* view, const MpsClockView* clk, int start_clock, const int* done_flag,
                    void* stream) {
  if (clk->model != MPS_MODEL_BSP || clk->own_count <= 0) return 0;
  auto st = static_cast<cudaStream_t>(stream);
  int grid = clock_grid(view, clk, 1);
  int dev = 0, sms = 148, per_sm = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  MPS_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, mps::fold_server_kernel, 256, 0));
  // every CTA must be resident (they wait on the workers together), and the servers of several tables
  // must fit side by side: at most one 256-thread CTA per SM each
  if (grid > sms * per_sm) grid = sms * per_sm;
  if (grid > sms) grid = sms;
  MpsTableView v = *view;
  MpsClockView c = *clk;
  void* args[] = {&v, &c, &start_clock, &done_flag};
  MPS_CHECK(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(mps::fold_server_kernel), dim3(grid), dim3(256), args,
                                        0, st));
  return 0;
}

int mps_barrier(const MpsClockView* clk, int epoch, void* stream) {
  mps::barrier_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(*clk, epoch);
  MPS_LAUNCH_CHECK();
  return 0;
}

int mps_straggle(long long ns, void* stream) {
  if (ns <= 0) return 0;
  mps::straggle_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(ns);
  MPS_LAUNCH_CHECK();
  return 0;
}

int mps_reset_pad(const MpsClockView* clk, int start_clock, void* stream) {
  mps::reset_pad_kernel<<<1, 256, 0, static_cast<cudaStream_t>(stream)>>>(*clk, start_clock);
  MPS_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
