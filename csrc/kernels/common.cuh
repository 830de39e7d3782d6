// common.cuh — device helpers shared by the sm_100a kernels: system-scope loads /
// stores / reductions on peer-mapped shards, owner lookup for range-partitioned tables,
// bounded spin-waits on the signal pad.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "minips_kernels.h"

#define MPS_CHECK(expr)                                   \
  do {                                                    \
    cudaError_t _e = (expr);                              \
    if (_e != cudaSuccess) return static_cast<int>(_e);   \
  } while (0)

#define MPS_LAUNCH_CHECK() MPS_CHECK(cudaGetLastError())

namespace mps {

// ---------------------------------------------------------------- memory-model helpers
// Weights of a shard may live on another GPU (NVLink peer mapping).  Peer lines are not
// kept coherent in the local L1, and other ranks add into them concurrently, so every
// access to shard memory is a strong .sys access (SASS: LDG/REDG ...STRONG.SYS).
__device__ __forceinline__ float ld_sys(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ double ld_sys(const double* p) {
  double v;
  asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_sys_v4(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ double2 ld_sys_v2(const double* p) {
  double2 v;
  asm volatile("ld.relaxed.sys.global.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_sys(float* p, float v) {
  asm volatile("red.relaxed.sys.global.add.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
__device__ __forceinline__ void red_sys(double* p, double v) {
  asm volatile("red.relaxed.sys.global.add.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
}
__device__ __forceinline__ void red_sys_v4(float* p, float4 v) {
  asm volatile("red.relaxed.sys.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x),
               "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ void red_gpu(float* p, float v) {
  asm volatile("red.relaxed.gpu.global.add.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_sys(const int* p) {
  int v;
  asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ int ld_relaxed_sys(const int* p) {
  int v;
  asm volatile("ld.relaxed.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(int* p, int v) {
  asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys(int* p, int v) {
  asm volatile("st.relaxed.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void fence_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }
__device__ __forceinline__ float4 ld_cg_v4(const float* p) {
  float4 v;
  asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ float ld_cg(const float* p) {
  float v;
  asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
// streaming (read-once) loads of the CSR arrays: keep them out of L1
__device__ __forceinline__ uint32_t ld_stream(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ float ld_stream(const float* p) {
  float v;
  asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}

// ---------------------------------------------------------------- range partition lookup
// Device analogue of RangePartitionManager::Slice (base/range_partition_manager.hpp:24-66):
// instead of cutting a sorted key array into per-server segments on the host, every key
// finds its owning shard itself.  Returns -1 for keys outside [begin[0], begin[S]) —
// the reference silently drops those.
__device__ __forceinline__ int owner_of(const MpsTableView& v, uint32_t key) {
  const int S = v.num_shards;
  if (key < v.begin[0] || key >= v.begin[S]) return -1;
  if (v.uniform_width) {
    int s = static_cast<int>((key - v.begin[0]) / v.uniform_width);
    return s < S ? s : S - 1;
  }
  int lo = 0, hi = S;  // invariant: begin[lo] <= key < begin[hi]
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (key >= v.begin[mid]) lo = mid; else hi = mid;
  }
  return lo;
}

// ---------------------------------------------------------------- pad / gate helpers
__device__ __forceinline__ bool pad_error(const MpsClockView& c) {
  return ld_relaxed_sys(c.pad[c.my_rank] + MPS_PAD_ERROR) != 0;
}

// min over the local replica of the progress table == ProgressTracker::min_clock_.
__device__ __forceinline__ int pad_min_clock(const MpsClockView& c) {
  const int* prog = c.pad[c.my_rank] + MPS_PAD_PROGRESS;
  int m = 0x7fffffff;
  for (int w = 0; w < c.num_workers; ++w) {
    int p = ld_acquire_sys(prog + w);
    m = p < m ? p : m;
  }
  return m;
}

// true when a Get of `worker` at clock `progress` may be served.
//   BSP : every shard folded up to `progress`   (bsp_model.cpp:45-56: reply iff progress == min)
//   SSP : progress <= min_clock + staleness      (ssp_model.cpp:58-85)
//   ASP : always                                 (asp_model.cpp:23-26)
__device__ __forceinline__ bool gate_open(const MpsClockView& c, int progress) {
  if (c.model == MPS_MODEL_ASP) return true;
  if (c.model == MPS_MODEL_SSP) return progress <= pad_min_clock(c) + c.staleness;
  const int* folded = c.pad[c.my_rank] + MPS_PAD_FOLDED;
  for (int s = 0; s < c.num_shards; ++s)
    if (ld_acquire_sys(folded + s) < progress) return false;
  return true;
}

// Single-thread bounded spin.  Returns false (and raises the sticky pad error) on timeout so
// a lost peer turns into an error code on the host instead of a hung GPU.
template <typename Pred>
__device__ __forceinline__ bool spin_until(const MpsClockView& c, Pred pred) {
  if (pred()) return true;
  const long long t0 = clock64();
  unsigned ns = 32;
  while (true) {
    if (pred()) return true;
    if (pad_error(c)) return false;
    if (clock64() - t0 > c.timeout_cycles) {
      st_relaxed_sys(c.pad[c.my_rank] + MPS_PAD_ERROR, 1);
      return false;
    }
    __nanosleep(ns);
    if (ns < 256) ns <<= 1;
  }
}

// Grid-wide barrier for cooperative (co-resident) launches: one arrival counter in global
// memory, monotonically increasing — the n-th barrier of a launch completes when the counter
// reaches n * gridDim.x.  The counter must be zero when the kernel starts.  Cheaper than
// cooperative_groups::grid_group::sync() (measured ~4 us on 296 CTAs) because only thread 0
// of each CTA fences, arrives and polls.
struct GridBarrier {
  int* ctr;
  int passed;
  __device__ __forceinline__ void sync() {
    __syncthreads();
    if (threadIdx.x == 0) {
      ++passed;
      const int target = passed * static_cast<int>(gridDim.x);
      __threadfence();
      atomicAdd(ctr, 1);
      int v;
      do {
        asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
      } while (v < target);
    }
    __syncthreads();
  }
};

// Whole-CTA gate: thread 0 spins, everyone else parks on the barrier.  The acquire in the
// spinning thread plus bar.sync orders the CTA's later shard loads after the flag.
__device__ __forceinline__ void cta_gate(const MpsClockView& c, int progress) {
  if (threadIdx.x == 0) spin_until(c, [&] { return gate_open(c, progress); });
  __syncthreads();
}

// ---------------------------------------------------------------- per-CTA completion counters
// "All G CTAs of rank r have finished phase X" without a last-CTA hop: every CTA publishes its own
// completion behind its own system fence (all CTAs in parallel) by adding 1 to slot `my_rank` of a
// counter on every rank; a reader that sees the full count has synchronised with every CTA's fence
// through the chain of reds.
// all `n` counters at `cnt` reached `target`?  (local spin on this rank's pad)
__device__ __forceinline__ bool all_reached(const int* cnt, int n, int target) {
  for (int r = 0; r < n; ++r)
    if (ld_acquire_sys(cnt + r) < target) return false;
  return true;
}
// this CTA's part is done: system fence, then +1 on slot `me` of the counter on every rank
__device__ __forceinline__ void cta_signal(const MpsClockView& c, int slot, bool multi) {
  if (multi) fence_sys(); else __threadfence();
  for (int i = 0; i < c.num_ranks; ++i) {
    const int r = (c.my_rank + i) % c.num_ranks;
    asm volatile("red.relaxed.sys.global.add.s32 [%0], %1;" ::"l"(c.pad[r] + slot + c.my_rank), "r"(1) : "memory");
  }
}

__device__ __forceinline__ uint32_t ld_sys_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace mps
