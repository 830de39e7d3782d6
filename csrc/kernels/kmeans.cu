// kmeans.cu — the non-GEMM half of a mini-batch k-means iteration.
//
// Reference (apps/kmeans/kmeans.cpp:238-273): for every point of the batch, sequentially,
// `lr = alpha / ++n_c; c -= lr * (c - x)`, then push new - old for all (K+1)*D keys and the K
// count deltas.  The sequential per-point update is replaced by its batched form:
//     S_c = points of the batch assigned to c,  m = |S_c|,
//     delta_c = alpha * (sum_{x in S_c} x - m * c) / (n_c + m),     delta_n_c = m
// computed from per-cluster partial sums, so the whole batch is assigned against the same pulled
// centres (what the tensor-core kernel needs) and the push is one dense slab.
#include <cuda_bf16.h>

#include "common.cuh"

namespace mps {

__device__ __forceinline__ void red_gpu_v4(float* p, float4 v) {
  asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}

// per-cluster partial sums / counts of a batch and its SSE: one warp per point, 4 dims per lane
__global__ void __launch_bounds__(256)
kmeans_accumulate_kernel(const __nv_bfloat16* __restrict__ x, const int32_t* __restrict__ assign,
                         const float* __restrict__ dist, int64_t n, int K, int D, float* __restrict__ sums,
                         float* __restrict__ counts, float* __restrict__ sse) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  float local_sse = 0.f;
  for (int64_t i = warp; i < n; i += nwarps) {
    const int a = assign[i];
    float xn = 0.f;
    for (int d0 = lane * 4; d0 < D; d0 += 128) {
      const uint2 raw = *reinterpret_cast<const uint2*>(x + i * D + d0);
      const float2 lo = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.x));
      const float2 hi = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.y));
      xn += lo.x * lo.x + lo.y * lo.y + hi.x * hi.x + hi.y * hi.y;
      if (a >= 0 && a < K) red_gpu_v4(sums + static_cast<int64_t>(a) * D + d0, make_float4(lo.x, lo.y, hi.x, hi.y));
    }
    xn = warp_sum(xn);
    if (lane == 0) {
      if (a >= 0 && a < K) red_gpu(counts + a, 1.f);
      local_sse += fmaxf(xn + dist[i], 0.f);
    }
  }
  if (sse) {
    local_sse = warp_sum(local_sse);
    if (lane == 0 && local_sse != 0.f) atomicAdd(sse, local_sse);
  }
}

__global__ void __launch_bounds__(256)
kmeans_delta_kernel(const float* __restrict__ centers, const float* __restrict__ counts_old,
                    const float* __restrict__ sums, const float* __restrict__ counts_batch, int K, int D, float alpha,
                    float* __restrict__ delta_c, float* __restrict__ delta_n) {
  const int64_t total = static_cast<int64_t>(K) * D;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int c = static_cast<int>(i / D);
    const float m = counts_batch[c];
    float d = 0.f;
    if (m > 0.f) d = alpha * (sums[i] - m * centers[i]) / (counts_old[c] + m);
    delta_c[i] = d;
    if (i % D == 0) delta_n[c] = m;
  }
}

// Centroid normalisation FUSED with the push of both tables (kmeans.cpp:247-274: the delta of every
// touched centre and its count, then Add + Add): one warp per centre row.  The old counts are read
// straight from the owning shard (the Get of the counts table, kmeans.cpp:224, under its consistency
// gate), the deltas go out as red.add into the owner's staging (BSP) or weights (SSP/ASP), and the
// batch accumulators are cleared for the next iteration, so no delta buffer and no separate
// pull / push / memset launch exists.
__global__ void __launch_bounds__(256)
kmeans_delta_push_kernel(const __grid_constant__ MpsTableView vc, const __grid_constant__ MpsTableView vn,
                         const __grid_constant__ MpsClockView cn, int gated, int progress,
                         const float* __restrict__ centers, float* __restrict__ sums,
                         float* __restrict__ counts_batch, int K, int D, float alpha) {
  if (gated) cta_gate(cn, progress);
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  for (int c = warp; c < K; c += nwarps) {
    const float m = counts_batch[c];
    if (m <= 0.f) continue;  // untouched centre: its sums row is still zero
    const uint32_t nkey = vn.begin[0] + static_cast<uint32_t>(c);
    const int sn = owner_of(vn, nkey);
    float n_old = 0.f;
    if (lane == 0) n_old = ld_sys(static_cast<const float*>(vn.weights[sn]) + (nkey - vn.begin[sn]));
    n_old = __shfl_sync(0xffffffffu, n_old, 0);
    const float scale = alpha / (n_old + m);
    for (int j = lane * 4; j < D; j += 128) {
      const int64_t f = static_cast<int64_t>(c) * D + j;
      const float4 s4 = *reinterpret_cast<const float4*>(sums + f);
      const float4 c4 = *reinterpret_cast<const float4*>(centers + f);
      float4 d;
      d.x = scale * (s4.x - m * c4.x);
      d.y = scale * (s4.y - m * c4.y);
      d.z = scale * (s4.z - m * c4.z);
      d.w = scale * (s4.w - m * c4.w);
      *reinterpret_cast<float4*>(sums + f) = make_float4(0.f, 0.f, 0.f, 0.f);
      // D % 4 == 0 and shard boundaries of the centre table fall on multiples of 4 values only when
      // the ranges do; route per element when the chunk straddles or is misaligned in its shard
      const uint32_t key0 = vc.begin[0] + static_cast<uint32_t>(f);
      const int sc = owner_of(vc, key0);
      const int64_t off = static_cast<int64_t>(key0) - vc.begin[sc];
      float* base = static_cast<float*>(vc.staging[sc] ? vc.staging[sc] : vc.weights[sc]);
      if (key0 + 4 <= vc.begin[sc + 1] && (off & 3) == 0) {
        if (d.x != 0.f || d.y != 0.f || d.z != 0.f || d.w != 0.f) red_sys_v4(base + off, d);
      } else {
        const float dv[4] = {d.x, d.y, d.z, d.w};
        for (int e = 0; e < 4; ++e) {
          const uint32_t key = key0 + e;
          const int se = owner_of(vc, key);
          float* be = static_cast<float*>(vc.staging[se] ? vc.staging[se] : vc.weights[se]);
          if (dv[e] != 0.f) red_sys(be + (key - vc.begin[se]), dv[e]);
        }
      }
    }
    __syncwarp();
    if (lane == 0) {
      float* nb = static_cast<float*>(vn.staging[sn] ? vn.staging[sn] : vn.weights[sn]);
      red_sys(nb + (nkey - vn.begin[sn]), m);
      counts_batch[c] = 0.f;
    }
  }
}

}  // namespace mps

extern "C" {

int mps_kmeans_accumulate(const void* x_bf16, const int32_t* assign, const float* dist, int64_t n, int K, int D,
                          float* sums, float* counts, float* sse, void* stream) {
  if (n <= 0) return 0;
  if (D % 4 != 0) return static_cast<int>(cudaErrorInvalidValue);
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  int64_t want = (n + 7) / 8;
  const int grid = static_cast<int>(want < sms * 8 ? want : sms * 8);
  mps::kmeans_accumulate_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x_bf16), assign, dist, n, K, D, sums, counts, sse);
  MPS_LAUNCH_CHECK();
  return 0;
}

int mps_kmeans_delta(const float* centers, const float* counts_old, const float* sums, const float* counts_batch,
                     int K, int D, float alpha, float* delta_c, float* delta_n, void* stream) {
  const int64_t total = static_cast<int64_t>(K) * D;
  const int grid = static_cast<int>((total + 255) / 256 < 1184 ? (total + 255) / 256 : 1184);
  mps::kmeans_delta_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(centers, counts_old, sums,
                                                                               counts_batch, K, D, alpha, delta_c,
                                                                               delta_n);
  MPS_LAUNCH_CHECK();
  return 0;
}

int mps_kmeans_delta_push(const MpsTableView* centers_view, const MpsTableView* counts_view,
                          const MpsClockView* counts_clk, int progress, const float* centers, float* sums,
                          float* counts_batch, int K, int D, float alpha, void* stream) {
  if (D % 4 != 0 || centers_view->dtype != MPS_F32 || counts_view->dtype != MPS_F32 ||
      centers_view->vals_per_key != 1 || counts_view->vals_per_key != 1)
    return static_cast<int>(cudaErrorInvalidValue);
  const int grid = (K + 7) / 8;  // 8 warps per CTA, one centre row per warp
  MpsClockView none{};
  mps::kmeans_delta_push_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      *centers_view, *counts_view, counts_clk ? *counts_clk : none, counts_clk != nullptr, progress, centers, sums,
      counts_batch, K, D, alpha);
  MPS_LAUNCH_CHECK();
  return 0;
}

int mps_kmeans_step(const MpsTableView* centers_view, const MpsClockView* centers_clk,
                    const MpsTableView* counts_view, const MpsClockView* counts_clk, int worker, int progress,
                    int fold, const void* x_bf16, int64_t n, int K, int D, float alpha,
                    void* centers_bf16_scratch, float* cnorm_scratch, float* centers_f32, int* sync_ctr,
                    int32_t* assign, float* dist, float* sums, float* counts_batch, float* sse, void* stream) {
  int rc = mps_kmeans_assign_tc(centers_view, centers_clk, progress, nullptr, x_bf16, n, K, D, centers_bf16_scratch,
                                cnorm_scratch, centers_f32, sync_ctr, assign, dist, 0, stream);
  if (rc) return rc;
  rc = mps_kmeans_accumulate(x_bf16, assign, dist, n, K, D, sums, counts_batch, sse, stream);
  if (rc) return rc;
  rc = mps_kmeans_delta_push(centers_view, counts_view, counts_clk, progress, centers_f32, sums, counts_batch, K, D,
                             alpha, stream);
  if (rc) return rc;
  rc = mps_clock(centers_view, centers_clk, worker, progress + 1, fold, stream);
  if (rc) return rc;
  return mps_clock(counts_view, counts_clk, worker, progress + 1, fold, stream);
}

}  // extern "C"
