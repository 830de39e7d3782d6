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

}  // extern "C"
