// kmeans_tc.cu — nearest-centre assignment on the 5th-gen tensor cores.
//
// Reference hot loop (apps/kmeans/kmeans_helper.hpp:45-66, called from apps/kmeans/kmeans.cpp:245):
// for every point, for every one of the K centres, copy the D-vector, subtract the sparse
// point and take the inner product — O(K*D) scalar work per point after a full-table pull of
// the (K+1)*D centroid table (kmeans.cpp:223).  Here:
//
//   argmin_c ||x - c||^2  =  argmin_c ( ||c||^2 - 2 x.c )
//
// is one persistent, warp-specialised kernel:
//   prologue  the centroid PULL fused into the B-operand fill: every CTA reads its slice of the
//             fp32 centroid table straight from the owning shards (peer ld over NVLink),
//             rounds to bf16 into the local B matrix and accumulates ||c||^2;
//   warp 0    TMA producer: A tiles (128 points x D, bf16) and B chunks (128 centres x D) into
//             128B-swizzled shared memory, mbarrier complete_tx;
//   warp 1    issues tcgen05.mma (M=128, N=128, K=16, bf16 -> fp32) into a ring of four
//             128-column TMEM accumulators; tcgen05.commit releases smem / signals the epilogue;
//   warps 2-5 epilogue: tcgen05.ld the accumulator, d = ||c||^2 - 2 acc, running arg-min with
//             the column index packed into the low mantissa bits (FFMA + LOP3 + FMNMX3 per
//             element, no distance matrix ever touches memory).
// A group of A_TILES point tiles stays resident in shared memory while the K/128 centre chunks
// stream past it, so B is re-read from L2 only once per 128*A_TILES points.
#include "tc_common.cuh"

namespace mps {
namespace kmtc {

constexpr int kTileM = 128;    // points per MMA tile
constexpr int kChunkN = 128;   // centres per MMA chunk
constexpr int kSlots = 4;      // TMEM accumulator ring (4 x 128 columns = all 512)
constexpr int kThreads = 320;  // warp 0 TMA, warp 1 MMA, warps 2..9 epilogue

using namespace mps::tc;  // PTX wrappers: tc_common.cuh

struct Params {
  MpsTableView centers;          // fp32 centroid table (K*D keys, range-partitioned)
  int has_view;                  // 0: centres_f32 below is a local dense [K, D] array instead
  const float* centers_f32;
  __nv_bfloat16* centers_bf16;   // scratch: [K, D] bf16(-2 c) | [K, 64] ||c||^2 split | [128, 64] ones
  float* cnorm;                  // [K] scratch: ||c||^2 of the rounded centres
  int* sync_ctr;                 // zeroed grid-barrier counter
  int32_t* assign;               // [n] out
  float* dist;                   // [n] out: ||c||^2 - 2 x.c of the chosen centre (add ||x||^2 for the SSE)
  int64_t n;
  int K;
  int K_valid;                   // centres [K_valid, K) are padding: zero coordinates, norm 2^127, never chosen
  int D_valid;                   // the table holds D_valid values per centre (row-major); columns beyond are zero
  MpsClockView clk;              // consistency gate of the centroid table (gated != 0)
  int gated;
  int progress;
  float* centers_f32_out;        // optional [K, D]: fp32 copy of the pulled centres (for the update step)
};

constexpr int kAugBytes = kChunkN * 128;  // one [128 x 64] bf16 swizzle-atom block

template <int kAtoms>  // D = 64 * kAtoms
struct Smem {
  static constexpr int D = 64 * kAtoms;
  static constexpr int kATiles = 3;
  static constexpr int kBStages = (D <= 64) ? 3 : 2;
  static constexpr int kTileBytes = kTileM * D * 2;
  static constexpr int kChunkBytes = kChunkN * D * 2;
  static constexpr int kStageBytes = kChunkBytes + kAugBytes;  // centre chunk + its ||c||^2 block
  static constexpr int kABytes = kATiles * kTileBytes;
  static constexpr int kOnesOff = kABytes;
  static constexpr int kBOff = kOnesOff + kAugBytes;
  static constexpr int kBarOff = kBOff + kBStages * kStageBytes;
  static constexpr int kNumBars = 2 * kATiles + 2 * kBStages + 2 * kSlots + 1;
  static constexpr int kExchOff = kBarOff + 8 * kNumBars + 16;  // epilogue half-merge buffer
  static constexpr int kEnd = kExchOff + kATiles * kTileM * 8;
  static int bytes() { return kEnd + 1024; }
};

// The epilogue is ALU-bound, so everything that can ride on the tensor pipe does: the centres
// are stored as bf16(-2c) (exact scaling) and ||c||^2 enters the accumulator through one extra
// K=16 MMA step  A' = [1 1 1 0 ...] (constant), B' = [hi mid lo 0 ...] with hi+mid+lo a
// three-way bf16 split of ||c||^2 (24 significant bits).  The accumulator then IS the distance
// term d = ||c||^2 - 2 x.c and the epilogue is LOP3 (pack index) + FMNMX3 per element.
template <int kAtoms>
__global__ void __launch_bounds__(kThreads, 1)
kmeans_assign_tc_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_c,
                        const __grid_constant__ CUtensorMap map_cn, const __grid_constant__ CUtensorMap map_ones,
                        const __grid_constant__ Params p) {
  using S = Smem<kAtoms>;
  constexpr int D = S::D;
  constexpr int kATiles = S::kATiles;
  constexpr int kBStages = S::kBStages;
  constexpr int kKSteps = D / 16;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sOnes = smem + S::kOnesOff;
  uint8_t* sB = smem + S::kBOff;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S::kBarOff);
  uint64_t* a_full = bars;
  uint64_t* a_empty = a_full + kATiles;
  uint64_t* b_full = a_empty + kATiles;
  uint64_t* b_empty = b_full + kBStages;
  uint64_t* t_full = b_empty + kBStages;
  uint64_t* t_empty = t_full + kSlots;
  uint64_t* ones_full = t_empty + kSlots;
  uint32_t* tmem_base_smem = reinterpret_cast<uint32_t*>(ones_full + 1);
  float* s_exch = reinterpret_cast<float*>(smem + S::kExchOff);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int K = p.K;
  const int n_chunks = K / kChunkN;
  __nv_bfloat16* g_cn = p.centers_bf16 + static_cast<int64_t>(K) * D;  // [K, 64]
  __nv_bfloat16* g_ones = g_cn + static_cast<int64_t>(K) * 64;         // [128, 64]

  // ---------------------------------------------------------- prologue: pull + scale/round + norms
  if (p.gated) cta_gate(p.clk, p.progress);
  {
    const int64_t gwarp = (static_cast<int64_t>(blockIdx.x) * kThreads + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * kThreads) >> 5;
    const int Kv = p.K_valid, Dv = p.D_valid;
    for (int64_t c = gwarp; c < K; c += nwarps) {
      float sq = 0.f;
      for (int d0 = lane * 4; d0 < D; d0 += 128) {
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        const int64_t flat = c * D + d0;  // position in the padded [K, D] operand
        if (c >= Kv) {
          // padding row: stays zero
        } else if (Dv != D) {
          // table rows are D_valid long: element-wise, zero beyond the row
          float t[4] = {0.f, 0.f, 0.f, 0.f};
          for (int j = 0; j < 4; ++j) {
            if (d0 + j >= Dv) break;
            const int64_t tf = c * Dv + d0 + j;
            if (p.has_view) {
              const uint32_t kj = p.centers.begin[0] + static_cast<uint32_t>(tf);
              const int sj = owner_of(p.centers, kj);
              t[j] = ld_sys(static_cast<const float*>(p.centers.weights[sj]) + (kj - p.centers.begin[sj]));
            } else {
              t[j] = p.centers_f32[tf];
            }
          }
          v = make_float4(t[0], t[1], t[2], t[3]);
        } else if (p.has_view) {
          const uint32_t key = p.centers.begin[0] + static_cast<uint32_t>(flat);
          const int s = owner_of(p.centers, key);
          const float* src = static_cast<const float*>(p.centers.weights[s]) + (key - p.centers.begin[s]);
          // shard boundaries fall on multiples of 4 keys whenever D % 4 == 0 and ranges are row aligned
          if (key + 3 < p.centers.begin[s + 1] && ((key - p.centers.begin[s]) & 3) == 0) {
            v = ld_sys_v4(src);
          } else {
            float t[4];
            for (int j = 0; j < 4; ++j) {
              const uint32_t kj = key + j;
              const int sj = owner_of(p.centers, kj);
              t[j] = ld_sys(static_cast<const float*>(p.centers.weights[sj]) + (kj - p.centers.begin[sj]));
            }
            v = make_float4(t[0], t[1], t[2], t[3]);
          }
        } else {
          v = *reinterpret_cast<const float4*>(p.centers_f32 + flat);
        }
        const __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
        const float2 flo = __bfloat1622float2(lo), fhi = __bfloat1622float2(hi);
        sq += flo.x * flo.x + flo.y * flo.y + fhi.x * fhi.x + fhi.y * fhi.y;
        const __nv_bfloat162 slo = __floats2bfloat162_rn(-2.f * flo.x, -2.f * flo.y);  // exact
        const __nv_bfloat162 shi = __floats2bfloat162_rn(-2.f * fhi.x, -2.f * fhi.y);
        uint2 packed;
        packed.x = *reinterpret_cast<const uint32_t*>(&slo);
        packed.y = *reinterpret_cast<const uint32_t*>(&shi);
        *reinterpret_cast<uint2*>(p.centers_bf16 + flat) = packed;
        if (p.centers_f32_out) *reinterpret_cast<float4*>(p.centers_f32_out + flat) = v;
      }
      sq = warp_sum(sq);
      if (c >= Kv) sq = 1.7014118346046923e38f;  // 2^127: exact in bf16, larger than any real distance term
      // ||c||^2 = hi + mid + lo in bf16, columns 0..2 of the augmentation block, zeros elsewhere
      const __nv_bfloat16 h = __float2bfloat16_rn(sq);
      const float r1 = sq - __bfloat162float(h);
      const __nv_bfloat16 m = __float2bfloat16_rn(r1);
      const __nv_bfloat16 l = __float2bfloat16_rn(r1 - __bfloat162float(m));
      const __nv_bfloat16 z = __float2bfloat16_rn(0.f);
      __nv_bfloat162 pair;
      pair.x = lane == 0 ? h : (lane == 1 ? l : z);
      pair.y = lane == 0 ? m : z;
      *reinterpret_cast<__nv_bfloat162*>(g_cn + c * 64 + lane * 2) = pair;
      if (lane == 0) p.cnorm[c] = sq;
    }
    if (blockIdx.x == 0) {
      const __nv_bfloat16 one = __float2bfloat16_rn(1.f), z = __float2bfloat16_rn(0.f);
      for (int i = threadIdx.x; i < kTileM * 64; i += kThreads) g_ones[i] = (i & 63) < 3 ? one : z;
    }
    // the operands are consumed by TMA (async proxy): order the generic-proxy writes before it
    fence_proxy_async();
    GridBarrier gb{p.sync_ctr, 0};
    gb.sync();
  }

  // ---------------------------------------------------------- barriers + TMEM
  if (threadIdx.x == 0) {
    for (int i = 0; i < kATiles; ++i) { mbar_init(a_full + i, 1); mbar_init(a_empty + i, 1); }
    for (int i = 0; i < kBStages; ++i) { mbar_init(b_full + i, 1); mbar_init(b_empty + i, 1); }
    for (int i = 0; i < kSlots; ++i) { mbar_init(t_full + i, 1); mbar_init(t_empty + i, 8); }
    mbar_init(ones_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_base_smem, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_base_smem;

  const int64_t group_rows = static_cast<int64_t>(kATiles) * kTileM;
  const int64_t n_groups = (p.n + group_rows - 1) / group_rows;

  if (warp == 0) {
    // ======================================================= TMA producer
    if (elect_one()) {
      mbar_expect_tx(ones_full, kAugBytes);
      tma_load_2d(sOnes, &map_ones, ones_full, 0, 0);
      uint32_t gi = 0, bi = 0;
      for (int64_t g = blockIdx.x; g < n_groups; g += gridDim.x, ++gi) {
        for (int c = 0; c < n_chunks; ++c, ++bi) {
          const uint32_t st = bi % kBStages;
          uint8_t* stage = sB + st * S::kStageBytes;
          mbar_wait(b_empty + st, ((bi / kBStages) & 1) ^ 1);
          mbar_expect_tx(b_full + st, S::kStageBytes);
          for (int a = 0; a < kAtoms; ++a)
            tma_load_2d(stage + a * (kChunkN * 128), &map_c, b_full + st, a * 64, c * kChunkN);
          tma_load_2d(stage + S::kChunkBytes, &map_cn, b_full + st, 0, c * kChunkN);
          if (c == 0) {
            for (int t = 0; t < kATiles; ++t) {
              mbar_wait(a_empty + t, (gi & 1) ^ 1);
              mbar_expect_tx(a_full + t, S::kTileBytes);
              const int row0 = static_cast<int>(g * group_rows + t * kTileM);
              for (int a = 0; a < kAtoms; ++a)
                tma_load_2d(sA + t * S::kTileBytes + a * (kTileM * 128), &map_x, a_full + t, a * 64, row0);
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ======================================================= MMA issuer
    constexpr uint32_t idesc = umma_idesc(kTileM, kChunkN);
    const uint64_t d_ones = umma_smem_desc(smem_u32(sOnes));
    mbar_wait(ones_full, 0);
    uint32_t gi = 0, bi = 0, ui = 0;
    for (int64_t g = blockIdx.x; g < n_groups; g += gridDim.x, ++gi) {
      for (int c = 0; c < n_chunks; ++c, ++bi) {
        const uint32_t st = bi % kBStages;
        const uint32_t stage = smem_u32(sB + st * S::kStageBytes);
        mbar_wait(b_full + st, (bi / kBStages) & 1);
        for (int t = 0; t < kATiles; ++t, ++ui) {
          if (c == 0) mbar_wait(a_full + t, gi & 1);
          const uint32_t slot = ui % kSlots;
          mbar_wait(t_empty + slot, ((ui / kSlots) & 1) ^ 1);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t d_tmem = tmem_base + slot * kChunkN;
            // ||c||^2 first (overwrites the slot), then the D/16 data steps accumulate -2 x.c
            umma_bf16(d_tmem, d_ones, umma_smem_desc(stage + S::kChunkBytes), idesc, 0u);
#pragma unroll
            for (int k = 0; k < kKSteps; ++k) {
              const int atom = k / 4, sub = k % 4;
              const uint64_t da = umma_smem_desc(smem_u32(sA + t * S::kTileBytes + atom * (kTileM * 128)) + sub * 32);
              const uint64_t db = umma_smem_desc(stage + atom * (kChunkN * 128) + sub * 32);
              umma_bf16(d_tmem, da, db, idesc, 1u);
            }
            umma_commit(t_full + slot);
            if (t == kATiles - 1) umma_commit(b_empty + st);
            if (c == n_chunks - 1) umma_commit(a_empty + t);
          }
          __syncwarp();
        }
      }
    }
  } else {
    // ======================================================= epilogue (warps 2..9)
    // Two warps per TMEM lane quarter (one per column half of every accumulator) so each SM
    // sub-partition has two epilogue warps to hide TMEM-load and ALU latency behind each other.
    const int quarter = warp & 3;        // TMEM lanes [32*quarter, 32*quarter + 32)
    const int half = (warp - 2) >> 2;    // accumulator columns [64*half, 64*half + 64)
    const float kInf = __int_as_float(0x7f7fffff);
    uint32_t ui = 0;
    for (int64_t g = blockIdx.x; g < n_groups; g += gridDim.x) {
      float best[kATiles];
      int bestc[kATiles];
#pragma unroll
      for (int t = 0; t < kATiles; ++t) { best[t] = kInf; bestc[t] = 0; }
      for (int c = 0; c < n_chunks; ++c) {
#pragma unroll
        for (int t = 0; t < kATiles; ++t, ++ui) {
          const uint32_t slot = ui % kSlots;
          mbar_wait(t_full + slot, (ui / kSlots) & 1);
          tc_fence_after();
          const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + slot * kChunkN + half * 64;
          uint32_t r0[32], r1[32];
          tmem_ld32(taddr, r0);
          tmem_ld32(taddr + 32, r1);
          tmem_ld_wait();
          // accumulator copied to registers: hand the slot back to the MMA warp right away
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(t_empty + slot);
          float m0 = kInf, m1 = kInf, m2 = kInf, m3 = kInf;  // independent chains
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const uint32_t col = static_cast<uint32_t>(half * 64 + j);
            m0 = fmin3(m0, __uint_as_float((r0[j + 0] & 0xFFFFFF80u) | (col + 0)),
                       __uint_as_float((r0[j + 1] & 0xFFFFFF80u) | (col + 1)));
            m1 = fmin3(m1, __uint_as_float((r0[j + 2] & 0xFFFFFF80u) | (col + 2)),
                       __uint_as_float((r0[j + 3] & 0xFFFFFF80u) | (col + 3)));
            m2 = fmin3(m2, __uint_as_float((r1[j + 0] & 0xFFFFFF80u) | (col + 32)),
                       __uint_as_float((r1[j + 1] & 0xFFFFFF80u) | (col + 33)));
            m3 = fmin3(m3, __uint_as_float((r1[j + 2] & 0xFFFFFF80u) | (col + 34)),
                       __uint_as_float((r1[j + 3] & 0xFFFFFF80u) | (col + 35)));
          }
          const float m = fminf(fmin3(m0, m1, m2), m3);
          if (m < best[t]) { best[t] = m; bestc[t] = c; }
        }
      }
      // merge the two column halves of every row, then the lower-half warps write the result
      float* ex_val = s_exch;
      int* ex_idx = reinterpret_cast<int*>(s_exch + kATiles * kTileM);
      if (half == 1) {
#pragma unroll
        for (int t = 0; t < kATiles; ++t) {
          ex_val[t * kTileM + quarter * 32 + lane] = best[t];
          ex_idx[t * kTileM + quarter * 32 + lane] = bestc[t];
        }
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (half == 0) {
#pragma unroll
        for (int t = 0; t < kATiles; ++t) {
          const float ov = ex_val[t * kTileM + quarter * 32 + lane];
          const int oc = ex_idx[t * kTileM + quarter * 32 + lane];
          if (ov < best[t]) { best[t] = ov; bestc[t] = oc; }
          const int64_t row = g * group_rows + t * kTileM + quarter * 32 + lane;
          if (row < p.n) {
            const uint32_t bits = __float_as_uint(best[t]);
            p.assign[row] = bestc[t] * kChunkN + static_cast<int>(bits & 127u);
            p.dist[row] = __uint_as_float(bits & 0xFFFFFF80u);
          }
        }
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");  // exchange buffer free for the next group
    }
  }

  // ---------------------------------------------------------- teardown
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

// Plain CUDA-core kernel of the same operation on the same bf16-rounded inputs: the
// differential reference for tests and the "recompiled legacy path" timing baseline.
__global__ void __launch_bounds__(256)
kmeans_assign_ref_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ c,
                         const float* __restrict__ cnorm, int64_t n, int K, int D, int32_t* __restrict__ assign,
                         float* __restrict__ dist) {
  const int64_t row = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (row >= n) return;
  float best = 3.4e38f;
  int bi = 0;
  for (int k = 0; k < K; ++k) {
    float acc = 0.f;
    for (int d = 0; d < D; ++d) acc = fmaf(__bfloat162float(x[row * D + d]),
        __bfloat162float(c[static_cast<int64_t>(k) * D + d]), acc);
    const float v = acc + cnorm[k];  // c is stored as bf16(-2 c)
    if (v < best) { best = v; bi = k; }
  }
  assign[row] = bi;
  dist[row] = best;
}

static int make_map(CUtensorMap* map, const void* base, int64_t rows, int D) {
  return make_map_bf16(map, base, rows, D, 128);
}

template <int kAtoms>
static int launch(const Params& p, const void* x, int max_ctas, cudaStream_t st) {
  using S = Smem<kAtoms>;
  CUtensorMap map_x, map_c, map_cn, map_ones;
  int rc = make_map(&map_x, x, p.n, S::D);
  if (rc) return rc;
  rc = make_map(&map_c, p.centers_bf16, p.K, S::D);
  if (rc) return rc;
  const __nv_bfloat16* cn = p.centers_bf16 + static_cast<int64_t>(p.K) * S::D;
  rc = make_map(&map_cn, cn, p.K, 64);
  if (rc) return rc;
  rc = make_map(&map_ones, cn + static_cast<int64_t>(p.K) * 64, kTileM, 64);
  if (rc) return rc;
  const int smem = S::bytes();
  auto kern = kmeans_assign_tc_kernel<kAtoms>;
  MPS_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int64_t group_rows = static_cast<int64_t>(S::kATiles) * kTileM;
  const int64_t n_groups = (p.n + group_rows - 1) / group_rows;
  int grid = static_cast<int>(n_groups < sms ? n_groups : sms);
  if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
  if (grid < 1) grid = 1;
  MPS_CHECK(cudaMemsetAsync(p.sync_ctr, 0, sizeof(int), st));
  Params pp = p;
  void* args[] = {&map_x, &map_c, &map_cn, &map_ones, &pp};
  MPS_CHECK(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(kern), dim3(grid), dim3(kThreads), args, smem, st));
  return 0;
}

}  // namespace kmtc
}  // namespace mps

extern "C" {

int mps_kmeans_assign_tc(const MpsTableView* centers_view, const MpsClockView* clk, int progress,
                         const float* centers_f32, const void* x_bf16, int64_t n, int K, int D, int K_valid,
                         int D_valid, void* centers_bf16_scratch, float* cnorm_scratch, float* centers_f32_out,
                         int* sync_ctr, int32_t* assign, float* dist, int max_ctas, void* stream) {
  if (n <= 0) return 0;
  if (K % mps::kmtc::kChunkN != 0 || K <= 0 || K > 8192) return static_cast<int>(cudaErrorInvalidValue);
  mps::kmtc::Params p{};
  if (centers_view) {
    p.centers = *centers_view;
    p.has_view = 1;
    if (centers_view->dtype != MPS_F32) return static_cast<int>(cudaErrorInvalidValue);
  } else {
    p.has_view = 0;
  }
  p.centers_f32 = centers_f32;
  p.centers_bf16 = static_cast<__nv_bfloat16*>(centers_bf16_scratch);
  p.cnorm = cnorm_scratch;
  p.sync_ctr = sync_ctr;
  p.assign = assign;
  p.dist = dist;
  p.n = n;
  p.K = K;
  p.K_valid = (K_valid > 0 && K_valid <= K) ? K_valid : K;
  p.D_valid = (D_valid > 0 && D_valid <= D) ? D_valid : D;
  p.gated = (clk != nullptr && centers_view != nullptr) ? 1 : 0;
  if (p.gated) p.clk = *clk;
  p.progress = progress;
  p.centers_f32_out = centers_f32_out;
  auto st = static_cast<cudaStream_t>(stream);
  if (D == 64) return mps::kmtc::launch<1>(p, x_bf16, max_ctas, st);
  if (D == 128) return mps::kmtc::launch<2>(p, x_bf16, max_ctas, st);
  return static_cast<int>(cudaErrorInvalidValue);
}

int mps_kmeans_assign_ref(const void* x_bf16, const void* centers_bf16, const float* cnorm, int64_t n, int K, int D,
                          int32_t* assign, float* dist, void* stream) {
  if (n <= 0) return 0;
  const int grid = static_cast<int>((n + 255) / 256);
  mps::kmtc::kmeans_assign_ref_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x_bf16), static_cast<const __nv_bfloat16*>(centers_bf16), cnorm, n, K, D,
      assign, dist);
  MPS_LAUNCH_CHECK();
  return 0;
}

/* 1 if an mbarrier wait of a tcgen05 kernel of this file expired since the last call (clears the flag);
 * synchronises with the device. */
int mps_kmeans_tc_error(void) { return mps::tc::tc_take_timeout_flag(); }

}  // extern "C"
