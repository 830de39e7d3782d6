// kmeans_runsum_bulk.cuh — run-sum of the centre-sorted batch with the row gathers done by the bulk
// copy engine (cp.async.bulk global -> shared, mbarrier completion) instead of register loads.
//
// The register version (kmeans_sorted_sum_kernel in kmeans.cu) is bound by the latency of random
// 256-byte row gathers: what it can keep in flight is limited by registers (8 rows x 16 B per lane).
// Here a warp owns a run of 64 sorted rows and puts ALL of them in flight at once: 64 bulk copies of one
// row each into its 16 KB slice of shared memory, four mbarriers (16 rows each) so summing starts as soon
// as the first quarter has landed.  4 warps per CTA = 64 KB of shared memory, 3 CTAs per SM = 192 KB of
// gathers in flight per SM.  D = 128 only (one row = 256 B = 8 B per lane).
//
// MEASURED (1x B200, 1 Mi points, K=1024): the whole k-means iteration takes 0.452 ms with this kernel against
// 0.378 ms with the register version, i.e. one bulk copy per 256-byte row costs more in the copy engine than
// it saves in latency.  Kept as an opt-in (MINIPS_RUNSUM=bulk) for differential testing; not the default.
#pragma once
#include "tc_common.cuh"

namespace mps {

constexpr int kBulkRun = 64;      // rows per warp-run (all in flight)
constexpr int kBulkStage = 16;    // rows per mbarrier
constexpr int kBulkWarps = 4;     // warps per CTA
constexpr int kBulkRowBytes = 256;
constexpr int kBulkSmem = kBulkWarps * kBulkRun * kBulkRowBytes + 64 * kBulkWarps;

__device__ __forceinline__ void bulk_row_g2s(void* smem_dst, const void* gsrc, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          tc::smem_u32(smem_dst)),
      "l"(gsrc), "r"(kBulkRowBytes), "r"(tc::smem_u32(bar))
      : "memory");
}

__global__ void __launch_bounds__(kBulkWarps * 32, 3)
kmeans_runsum_bulk_kernel(const __nv_bfloat16* __restrict__ x, const int32_t* __restrict__ perm,
                          const int* __restrict__ base, const float* __restrict__ dist, int K,
                          float* __restrict__ sums, float* __restrict__ sse) {
  constexpr int D = 128;
  extern __shared__ __align__(128) unsigned char smem[];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  unsigned char* rows = smem + w * kBulkRun * kBulkRowBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kBulkWarps * kBulkRun * kBulkRowBytes) + w * 8;
  if (lane < 4) tc::mbar_init(&bars[lane], 1);
  tc::fence_barrier_init();
  tc::fence_proxy_async();
  __syncwarp();
  const int total = base[K];
  const int64_t gw = static_cast<int64_t>(blockIdx.x) * kBulkWarps + w;
  const int64_t nwarps = static_cast<int64_t>(gridDim.x) * kBulkWarps;
  uint32_t phase = 0;
  float local_sse = 0.f;
  for (int64_t run = gw; run * kBulkRun < total; run += nwarps) {
    const int r0 = static_cast<int>(run * kBulkRun);
    const int r1 = r0 + kBulkRun < total ? r0 + kBulkRun : total;
    const int i0 = r0 + lane < r1 ? perm[r0 + lane] : -1;
    const int i1 = r0 + 32 + lane < r1 ? perm[r0 + 32 + lane] : -1;
    if (lane < 4) {
      int valid = r1 - (r0 + kBulkStage * lane);
      valid = valid < 0 ? 0 : (valid > kBulkStage ? kBulkStage : valid);
      tc::mbar_expect_tx(&bars[lane], static_cast<uint32_t>(valid) * kBulkRowBytes);
    }
    __syncwarp();
    if (i0 >= 0) bulk_row_g2s(rows + lane * kBulkRowBytes, x + static_cast<int64_t>(i0) * D, &bars[lane >> 4]);
    if (i1 >= 0)
      bulk_row_g2s(rows + (32 + lane) * kBulkRowBytes, x + static_cast<int64_t>(i1) * D, &bars[2 + (lane >> 4)]);
    const float d0 = i0 >= 0 ? dist[i0] : 0.f;
    const float d1 = i1 >= 0 ? dist[i1] : 0.f;
    // centre of the first row: largest c with base[c] <= r0
    int lo = 0, hi = K;
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (base[mid] <= r0) lo = mid; else hi = mid;
    }
    int c = lo;
    int next = base[c + 1];
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int s = 0; s < kBulkRun / kBulkStage; ++s) {
      if (r0 + s * kBulkStage >= r1) break;
      tc::mbar_wait(&bars[s], phase);
      for (int j = 0; j < kBulkStage; ++j) {
        const int row = s * kBulkStage + j;
        const int pos = r0 + row;
        if (pos >= r1) break;
        const uint2 raw = *reinterpret_cast<const uint2*>(rows + row * kBulkRowBytes + lane * 8);
        if (pos >= next) {  // the run crosses into another centre
          red_gpu_v4(sums + static_cast<int64_t>(c) * D + lane * 4, acc);
          acc = make_float4(0.f, 0.f, 0.f, 0.f);
          do { ++c; next = base[c + 1]; } while (pos >= next);
        }
        const float2 p0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.x));
        const float2 p1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.y));
        acc.x += p0.x; acc.y += p0.y; acc.z += p1.x; acc.w += p1.y;
        const float xn = warp_sum(p0.x * p0.x + p0.y * p0.y + p1.x * p1.x + p1.y * p1.y);
        const float d = __shfl_sync(0xffffffffu, row < 32 ? d0 : d1, row & 31);
        if (lane == 0) local_sse += fmaxf(xn + d, 0.f);
      }
    }
    red_gpu_v4(sums + static_cast<int64_t>(c) * D + lane * 4, acc);
    phase ^= 1u;
    __syncwarp();            // every lane is done reading the slice ...
    tc::fence_proxy_async(); // ... before the copy engine overwrites it for the next run
  }
  if (sse && lane == 0 && local_sse != 0.f) atomicAdd(sse, local_sse);
}

}  // namespace mps
