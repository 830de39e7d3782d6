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

#include <cstdlib>

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

// ---- sorted accumulation -----------------------------------------------------------------
// The kernel above issues one red.add.v4 per point and lane (32 M L2 atomics for a 1 M-point batch),
// which costs as much as the distance GEMM.  The sorted path groups the batch by centre first
// (counting sort of the row indices: histogram -> scan -> scatter), then every warp walks a run of
// 64 consecutive sorted rows with the partial sum of the current centre in registers and issues one
// red.add.v4 per lane only when the centre changes: HBM-bound row gathers instead of atomics.
__global__ void __launch_bounds__(256)
kmeans_hist_kernel(const int32_t* __restrict__ assign, int64_t n, int K, int* __restrict__ hist) {
  extern __shared__ int sh[];
  for (int c = threadIdx.x; c < K; c += blockDim.x) sh[c] = 0;
  __syncthreads();
  const int64_t chunk = (n + gridDim.x - 1) / gridDim.x;
  const int64_t lo = chunk * blockIdx.x, hi = (lo + chunk < n) ? lo + chunk : n;
  for (int64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
    const int a = assign[i];
    if (a >= 0 && a < K) atomicAdd(&sh[a], 1);
  }
  __syncthreads();
  for (int c = threadIdx.x; c < K; c += blockDim.x)
    if (sh[c]) atomicAdd(&hist[c], sh[c]);
}

// one CTA: base[c] = exclusive prefix of hist, cursor = base, counts[c] += hist[c], hist cleared
__global__ void __launch_bounds__(1024)
kmeans_scan_kernel(int* __restrict__ hist, int K, int* __restrict__ base, int* __restrict__ cursor,
                   float* __restrict__ counts) {
  __shared__ int warp_tot[32];
  __shared__ int carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int c0 = 0; c0 < K; c0 += blockDim.x) {
    const int c = c0 + threadIdx.x;
    const int h = c < K ? hist[c] : 0;
    int incl = h;
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) warp_tot[w] = incl;
    __syncthreads();
    if (w == 0) {
      int t = warp_tot[lane];
      for (int o = 1; o < 32; o <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, t, o);
        if (lane >= o) t += u;
      }
      warp_tot[lane] = t;  // inclusive totals of the warps
    }
    __syncthreads();
    const int excl = carry + (w ? warp_tot[w - 1] : 0) + incl - h;
    if (c < K) {
      base[c] = excl;
      cursor[c] = excl;
      if (h) counts[c] += static_cast<float>(h);
      hist[c] = 0;
    }
    __syncthreads();
    if (threadIdx.x == 0) carry += warp_tot[31];
    __syncthreads();
  }
  if (threadIdx.x == 0) base[K] = carry;
}

__global__ void __launch_bounds__(256)
kmeans_scatter_kernel(const int32_t* __restrict__ assign, int64_t n, int K, int* __restrict__ cursor,
                      int32_t* __restrict__ perm) {
  extern __shared__ int sh[];
  int* cnt = sh;        // rows of this CTA's chunk per centre, then the fill counter
  int* start = sh + K;  // first output slot reserved for (this CTA, centre)
  for (int c = threadIdx.x; c < K; c += blockDim.x) cnt[c] = 0;
  __syncthreads();
  const int64_t chunk = (n + gridDim.x - 1) / gridDim.x;
  const int64_t lo = chunk * blockIdx.x, hi = (lo + chunk < n) ? lo + chunk : n;
  for (int64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
    const int a = assign[i];
    if (a >= 0 && a < K) atomicAdd(&cnt[a], 1);
  }
  __syncthreads();
  for (int c = threadIdx.x; c < K; c += blockDim.x) {
    const int m = cnt[c];
    if (m) start[c] = atomicAdd(&cursor[c], m);
    cnt[c] = 0;
  }
  __syncthreads();
  for (int64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
    const int a = assign[i];
    if (a >= 0 && a < K) perm[start[a] + atomicAdd(&cnt[a], 1)] = static_cast<int32_t>(i);
  }
}

}  // namespace mps
#include "kmeans_runsum_bulk.cuh"
namespace mps {

constexpr int kSortedRun = 64;  // sorted rows per unit

// One "unit" = D/8 lanes (16 B = 8 bf16 per lane, so a unit covers one row per load); a warp holds
// 32 / (D/8) independent units.  Every unit owns a run of kSortedRun sorted rows and keeps 8 row loads
// in flight.  All loops have warp-uniform trip counts (rows past the end are predicated off) so the
// shuffles stay convergent while units flush at different positions.
template <int D>
__global__ void __launch_bounds__(256, 3)
kmeans_sorted_sum_kernel(const __nv_bfloat16* __restrict__ x, const int32_t* __restrict__ perm,
                         const int* __restrict__ base, const float* __restrict__ dist, int K,
                         float* __restrict__ sums, float* __restrict__ sse) {
  static_assert(D == 64 || D == 128, "8 bf16 per lane");
  constexpr int LPR = D / 8;          // lanes per row
  constexpr int UPW = 32 / LPR;       // units per warp
  constexpr int U = 8;                // row loads in flight per unit
  static_assert(LPR % U == 0, "a perm block is consumed in whole groups of U rows");
  const int lane = threadIdx.x & 31;
  const int sub = lane & (LPR - 1);   // lane within the unit
  const int ubase = lane & ~(LPR - 1);
  const int total = base[K];
  const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  float local_sse = 0.f;
  for (int64_t w0 = warp; w0 * UPW * kSortedRun < total; w0 += nwarps) {
    const int64_t unit = w0 * UPW + lane / LPR;
    const int r0 = static_cast<int>(unit * kSortedRun < total ? unit * kSortedRun : total);
    const int r1 = r0 + kSortedRun < total ? r0 + kSortedRun : total;
    int c = 0, next = 0x7fffffff;
    if (r0 < r1) {  // centre of the first row: largest c with base[c] <= r0
      int lo = 0, hi = K;
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (base[mid] <= r0) lo = mid; else hi = mid;
      }
      c = lo;
      next = base[c + 1];
    }
    float acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.f;
    for (int b0 = 0; b0 < kSortedRun; b0 += LPR) {
      const int pl = r0 + b0 + sub;
      const int mine = pl < r1 ? perm[pl] : -1;
      const float my_dist = mine >= 0 ? dist[mine] : 0.f;
#pragma unroll
      for (int j0 = 0; j0 < LPR; j0 += U) {
        uint4 raw[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int idx = __shfl_sync(0xffffffffu, mine, ubase + j0 + u);
          raw[u] = make_uint4(0u, 0u, 0u, 0u);
          if (idx >= 0) raw[u] = *reinterpret_cast<const uint4*>(x + static_cast<int64_t>(idx) * D + sub * 8);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int pos = r0 + b0 + j0 + u;
          const bool valid = pos < r1;  // rows past the end of the unit's run were loaded as zeros
          const float2 p0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw[u].x));
          const float2 p1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw[u].y));
          const float2 p2 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw[u].z));
          const float2 p3 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw[u].w));
          if (valid && pos >= next) {  // this unit's run crosses into another centre
            float* dst = sums + static_cast<int64_t>(c) * D + sub * 8;
            red_gpu_v4(dst, make_float4(acc[0], acc[1], acc[2], acc[3]));
            red_gpu_v4(dst + 4, make_float4(acc[4], acc[5], acc[6], acc[7]));
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[e] = 0.f;
            do { ++c; next = base[c + 1]; } while (pos >= next);
          }
          acc[0] += p0.x; acc[1] += p0.y; acc[2] += p1.x; acc[3] += p1.y;
          acc[4] += p2.x; acc[5] += p2.y; acc[6] += p3.x; acc[7] += p3.y;
          float xn = p0.x * p0.x + p0.y * p0.y + p1.x * p1.x + p1.y * p1.y + p2.x * p2.x + p2.y * p2.y +
                     p3.x * p3.x + p3.y * p3.y;
#pragma unroll
          for (int o = LPR / 2; o > 0; o >>= 1) xn += __shfl_xor_sync(0xffffffffu, xn, o);
          const float d = __shfl_sync(0xffffffffu, my_dist, ubase + j0 + u);
          if (sub == 0 && valid) local_sse += fmaxf(xn + d, 0.f);
        }
      }
    }
    if (r0 < r1) {
      float* dst = sums + static_cast<int64_t>(c) * D + sub * 8;
      red_gpu_v4(dst, make_float4(acc[0], acc[1], acc[2], acc[3]));
      red_gpu_v4(dst + 4, make_float4(acc[4], acc[5], acc[6], acc[7]));
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
                         float* __restrict__ counts_batch, int K, int D, int Dv, float alpha) {
  // K = number of real centres; `centers` / `sums` are [*, D] (padded operand layout), the table rows
  // are Dv values long
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
      const uint32_t key0 = vc.begin[0] + static_cast<uint32_t>(static_cast<int64_t>(c) * Dv + j);
      const int sc = owner_of(vc, key0);
      const int64_t off = static_cast<int64_t>(key0) - vc.begin[sc];
      float* base = static_cast<float*>(vc.staging[sc] ? vc.staging[sc] : vc.weights[sc]);
      if (Dv == D && key0 + 4 <= vc.begin[sc + 1] && (off & 3) == 0) {
        if (d.x != 0.f || d.y != 0.f || d.z != 0.f || d.w != 0.f) red_sys_v4(base + off, d);
      } else {
        const float dv[4] = {d.x, d.y, d.z, d.w};
        for (int e = 0; e < 4; ++e) {
          if (j + e >= Dv) break;  // padding columns carry no key
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

int mps_kmeans_accumulate_sorted(const void* x_bf16, const int32_t* assign, const float* dist, int64_t n, int K,
                                 int D, float* sums, float* counts, float* sse, int* sort_scratch, void* stream) {
  if (n <= 0) return 0;
  if ((D != 64 && D != 128) || K > 4096 || n > (int64_t(1) << 30) || !sort_scratch)
    return mps_kmeans_accumulate(x_bf16, assign, dist, n, K, D, sums, counts, sse, stream);
  auto st = static_cast<cudaStream_t>(stream);
  int* hist = sort_scratch;          // [K]   zero on entry, left zero
  int* base = hist + K;              // [K+1]
  int* cursor = base + K + 1;        // [K]
  int32_t* perm = cursor + K;        // [n]
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  }
  int64_t want = (n + 4095) / 4096;  // >= 16 rows per thread so the per-CTA histogram pays
  const int grid = static_cast<int>(want < sms * 4 ? want : sms * 4);
  mps::kmeans_hist_kernel<<<grid, 256, K * sizeof(int), st>>>(assign, n, K, hist);
  mps::kmeans_scan_kernel<<<1, 1024, 0, st>>>(hist, K, base, cursor, counts);
  mps::kmeans_scatter_kernel<<<grid, 256, 2 * K * sizeof(int), st>>>(assign, n, K, cursor, perm);
  const int64_t runs = (n + mps::kSortedRun - 1) / mps::kSortedRun;
  const int upw = 32 / (D / 8);                       // units per warp
  const int64_t ctas = (runs + 8 * upw - 1) / (8 * upw);
  const int g2 = static_cast<int>(ctas < sms * 8 ? ctas : sms * 8);
  // MINIPS_RUNSUM=bulk selects the copy-engine version (kmeans_runsum_bulk.cuh); measured slower, see there
  static int use_bulk = -1;
  if (use_bulk < 0) {
    const char* e = getenv("MINIPS_RUNSUM");
    use_bulk = (e && e[0] == 'b') ? 1 : 0;
    if (use_bulk && cudaFuncSetAttribute(mps::kmeans_runsum_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         mps::kBulkSmem) != cudaSuccess) {
      cudaGetLastError();
      use_bulk = 0;
    }
  }
  if (D == 128 && use_bulk) {
    const int64_t bruns = (n + mps::kBulkRun - 1) / mps::kBulkRun;
    const int64_t bctas = (bruns + mps::kBulkWarps - 1) / mps::kBulkWarps;
    const int bg = static_cast<int>(bctas < sms * 3 ? bctas : sms * 3);
    mps::kmeans_runsum_bulk_kernel<<<bg, mps::kBulkWarps * 32, mps::kBulkSmem, st>>>(
        static_cast<const __nv_bfloat16*>(x_bf16), perm, base, dist, K, sums, sse);
  } else if (D == 128)
    mps::kmeans_sorted_sum_kernel<128><<<g2, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(x_bf16), perm, base, dist,
                                                          K, sums, sse);
  else
    mps::kmeans_sorted_sum_kernel<64><<<g2, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(x_bf16), perm, base, dist,
                                                         K, sums, sse);
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
                          float* counts_batch, int K, int D, int K_valid, int D_valid, float alpha, void* stream) {
  const int Kv = (K_valid > 0 && K_valid <= K) ? K_valid : K;
  const int Dv = (D_valid > 0 && D_valid <= D) ? D_valid : D;
  if (D % 4 != 0 || centers_view->dtype != MPS_F32 || counts_view->dtype != MPS_F32 ||
      centers_view->vals_per_key != 1 || counts_view->vals_per_key != 1)
    return static_cast<int>(cudaErrorInvalidValue);
  const int grid = (Kv + 7) / 8;  // 8 warps per CTA, one centre row per warp
  MpsClockView none{};
  mps::kmeans_delta_push_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      *centers_view, *counts_view, counts_clk ? *counts_clk : none, counts_clk != nullptr, progress, centers, sums,
      counts_batch, Kv, D, Dv, alpha);
  MPS_LAUNCH_CHECK();
  return 0;
}

int mps_kmeans_step(const MpsTableView* centers_view, const MpsClockView* centers_clk,
                    const MpsTableView* counts_view, const MpsClockView* counts_clk, int worker, int progress,
                    int fold, const void* x_bf16, int64_t n, int K, int D, int K_valid, int D_valid, float alpha,
                    void* centers_bf16_scratch, float* cnorm_scratch, float* centers_f32, int* sync_ctr,
                    int32_t* assign, float* dist, float* sums, float* counts_batch, float* sse, int* sort_scratch,
                    void* stream) {
  int rc = mps_kmeans_assign_tc(centers_view, centers_clk, progress, nullptr, x_bf16, n, K, D, K_valid, D_valid,
                                centers_bf16_scratch,
                                cnorm_scratch, centers_f32, sync_ctr, assign, dist, 0, stream);
  if (rc) return rc;
  rc = mps_kmeans_accumulate_sorted(x_bf16, assign, dist, n, K, D, sums, counts_batch, sse, sort_scratch, stream);
  if (rc) return rc;
  rc = mps_kmeans_delta_push(centers_view, counts_view, counts_clk, progress, centers_f32, sums, counts_batch, K, D,
                             K_valid, D_valid, alpha, stream);
  if (rc) return rc;
  return mps_clock2(centers_view, centers_clk, counts_view, counts_clk, worker, progress + 1, fold, stream);
}

}  // extern "C"
