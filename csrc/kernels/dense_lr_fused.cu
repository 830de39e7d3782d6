// dense_lr_fused.cu — one dense logistic-regression iteration that reads X ONCE.
//
// dense_lr_tc.cu computes the logits on the tensor cores and then walks X a second time on the CUDA
// cores for the gradient (grad = coef^T X): twice the algorithmic bytes of a bandwidth-bound step.
// Here a 64-row tile of X stays in shared memory between the two products:
//
//   warp 0    TMA producer: 64-row x 64-column blocks of X (bf16, 128B swizzle) into a ring that holds one
//             tile plus a few blocks of the next;
//   warp 1    issues, for every tile,
//               forward   D1[rows, 16]  += X_blk[rows, 64k] . w_blk[16, 64k]^T     (A K-major, the usual way)
//               gradient  D2[128 features, 16] += X_blk^T[128 features, 64 rows] . coef[16, 64 rows]^T
//             the SAME shared-memory blocks serve as the MN-major ("transposed") A operand of the second
//             product, so the features of X land on the 128 TMEM lanes and the gradient of the whole
//             kernel accumulates in TMEM (16 columns per 128 features, never spilled per tile);
//   warps 2-5 epilogue: logits out of TMEM, sigmoid / logloss / coef = alpha (y - p); coef goes back
//             into shared memory as the B operand of the gradient product (as hi + lo bf16 rows, so the
//             product sees 16 mantissa bits); at the very end the gradient leaves TMEM with one
//             red.add per feature and CTA.
// The forward product runs with M = 128 on a 64-row tile (the upper 64 accumulator lanes read the next
// block's bytes and are ignored): tcgen05 M = 64 has a different TMEM layout and the tensor pipe has
// the headroom.  0.66 ms per iteration at 1 Mi x 1024 (two passes: 0.75 ms; dense_lr_stream.cu: 0.64 ms).
// Reference op: apps/logistic_regression/lr_example.cpp:293-312 for dense rows.
//
// Limits: D a multiple of 8, roundup(D, 128) <= 1024 (tile + lookahead must fit in shared memory);
// anything else takes mps_dense_lr_forward_tc + mps_dense_lr_grad.
#include "tc_common.cuh"

namespace mps {
namespace dlrf {

using namespace mps::tc;  // NOLINT

constexpr int kRows = 64;                    // rows of X per tile
constexpr int kBlk = kRows * 128;            // one [64 rows x 64 cols] bf16 block = 8 KiB
constexpr int kN = 16;                       // UMMA N of both products
constexpr int kWBlk = kN * 128;              // w block [16 x 64] = 2 KiB
constexpr int kCoefBytes = kN * 128;         // coef operand [16 x 64 samples] = 2 KiB
constexpr int kThreads = 192;                // warp 0 TMA, warp 1 MMA, warps 2..5 epilogue
constexpr int kMaxBlocks = 16;               // blocks per tile: Dp <= 1024
constexpr int kMaxRing = 24;                 // ring slots
constexpr int kAcc = 8;                      // independent forward accumulators (see the MMA issuer)
constexpr int kSlack = 128 / kRows - 1;      // blocks a forward MMA (M = 128) reads past its own
constexpr int kTmemCols = 512;               // 2 x 8 x 16 forward + up to 8 x 16 gradient

struct Params {
  MpsTableView weights;   // fp32 table of D keys
  int has_view;
  const float* w_f32;     // local dense weights when has_view == 0
  MpsClockView clk;
  int gated, progress;
  __nv_bfloat16* w16;     // scratch [16, Dp] bf16, row 0 = w
  float* w_f32_out;       // optional [D]
  int* sync_ctr;
  const float* label;     // [n] in {0,1}
  float alpha;
  float* grad;            // [D] += sum_r coef[r] x[r, :]
  float* loss_out;        // [2]: += logloss sum, += #correct
  int64_t n;
  int D, Dp;              // Dp = D rounded up to 128
  int ring;               // ring slots (blocks of X resident at a time)
};

// MN-major, 128B-swizzled operand (the K-major TMA block read the other way): 64 MN elements are
// contiguous (128 B), the next 64 MN elements are LBO away, 8 K rows are 128 B apart and the next 8 K
// rows SBO away.
__device__ __forceinline__ uint64_t umma_smem_desc_mn(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFF) >> 4);
  d |= static_cast<uint64_t>((lbo >> 4) & 0x3FFF) << 16;
  d |= static_cast<uint64_t>((sbo >> 4) & 0x3FFF) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}

__global__ void __launch_bounds__(kThreads, 1)
dense_lr_fused_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w,
                      const __grid_constant__ Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int R = p.ring;
  const int nb = p.Dp / 64;    // blocks per tile
  const int nfb = p.Dp / 128;  // 128-feature groups
  uint8_t* sX = smem;                                        // R blocks + slack (M = 128 over-read)
  uint8_t* sW = sX + static_cast<size_t>(R + kSlack) * kBlk; // nb w blocks
  uint8_t* sC = sW + static_cast<size_t>(nb) * kWBlk;   // 2 coef operands
  uint64_t* bars = reinterpret_cast<uint64_t*>(sC + 2 * kCoefBytes);
  uint64_t* x_full = bars;                 // [R]
  uint64_t* x_empty = x_full + kMaxRing;   // [R]
  uint64_t* w_full = x_empty + kMaxRing;   // [1]
  uint64_t* f_full = w_full + 1;           // [2] forward accumulator ready
  uint64_t* f_empty = f_full + 2;          // [2] forward accumulator drained
  uint64_t* c_full = f_empty + 2;          // [2] coef operand written
  uint64_t* c_empty = c_full + 2;          // [2] gradient MMAs of that tile retired
  uint64_t* g_full = c_empty + 2;          // [1] all gradient MMAs retired
  uint32_t* tmem_base_smem = reinterpret_cast<uint32_t*>(g_full + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // ---- prologue: pull w, round to bf16 into row 0 of the [16, Dp] operand (other rows stay zero)
  if (p.gated) cta_gate(p.clk, p.progress);
  {
    const int64_t tid = static_cast<int64_t>(blockIdx.x) * kThreads + threadIdx.x;
    const int64_t nth = static_cast<int64_t>(gridDim.x) * kThreads;
    for (int64_t i = tid; i < static_cast<int64_t>(kN) * p.Dp; i += nth) {
      const int64_t d = i % p.Dp;
      float v = 0.f;
      if (i < p.Dp && d < p.D) {
        if (p.has_view) {
          const uint32_t key = p.weights.begin[0] + static_cast<uint32_t>(d);
          const int s = owner_of(p.weights, key);
          v = ld_sys(static_cast<const float*>(p.weights.weights[s]) + (key - p.weights.begin[s]));
        } else {
          v = p.w_f32[d];
        }
        if (p.w_f32_out) p.w_f32_out[d] = v;
      }
      p.w16[i] = __float2bfloat16_rn(v);
    }
    fence_proxy_async();
    GridBarrier gb{p.sync_ctr, 0};
    gb.sync();
  }

  // coef operands start as zeros (rows 2..15 are never written)
  for (int i = threadIdx.x; i < 2 * kCoefBytes / 4; i += kThreads) reinterpret_cast<uint32_t*>(sC)[i] = 0u;
  if (threadIdx.x == 0) {
    for (int i = 0; i < R; ++i) { mbar_init(x_full + i, 1); mbar_init(x_empty + i, 1); }
    mbar_init(w_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(f_full + i, 1);
      mbar_init(f_empty + i, kRows / 32);
      mbar_init(c_full + i, kRows / 32);
      mbar_init(c_empty + i, 1);
    }
    mbar_init(g_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_base_smem, kTmemCols);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_base_smem;
  const uint32_t tmem_grad = tmem_base + 2 * kAcc * kN;
  const int64_t n_tiles = (p.n + kRows - 1) / kRows;
  const int64_t my_tiles = n_tiles > blockIdx.x ? (n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;

  if (warp == 0) {
    // ======================================================= TMA producer
    if (elect_one()) {
      mbar_expect_tx(w_full, nb * kWBlk);
      for (int b = 0; b < nb; ++b) tma_load_2d(sW + b * kWBlk, &map_w, w_full, b * 64, 0);
      // ring position kept incrementally: a 64-bit divide per block costs more than the block's MMAs
      uint32_t s = 0, ph = 0;
      for (int64_t i = 0; i < my_tiles; ++i) {
        const int64_t tile = blockIdx.x + i * gridDim.x;
        for (int b = 0; b < nb; ++b) {
          mbar_wait(x_empty + s, ph ^ 1);
          mbar_expect_tx(x_full + s, kBlk);
          tma_load_2d(sX + s * kBlk, &map_x, x_full + s, b * 64, static_cast<int>(tile * kRows));
          if (++s == static_cast<uint32_t>(R)) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ======================================================= MMA issuer
    constexpr uint32_t idesc_fwd = umma_idesc(128, kN);
    constexpr uint32_t idesc_grad = umma_idesc(128, kN) | (1u << 15);  // A is MN-major
    const int look = R - nb;  // blocks of the NEXT tile that fit beside a resident tile
    mbar_wait(w_full, 0);
    tc_fence_after();
    // forward MMAs of block b of local tile i (blocks are consumed in arrival order)
    uint32_t fslot = 0, fph = 0;  // ring slot / phase of the next forward block (issued in stream order)
    auto forward_block = [&](int64_t i, int b) {
      const uint32_t s = fslot;
      const uint32_t fs = static_cast<uint32_t>(i & 1);
      if (b == 0) {
        mbar_wait(f_empty + fs, ((i >> 1) & 1) ^ 1);
        tc_fence_after();
      }
      mbar_wait(x_full + s, fph);
      tc_fence_after();
      if (++fslot == static_cast<uint32_t>(R)) { fslot = 0; fph ^= 1; }
      if (elect_one()) {
        const uint32_t xa = smem_u32(sX + s * kBlk), wa = smem_u32(sW + b * kWBlk);
        // eight accumulators in rotation: consecutive MMAs never accumulate onto each other, the epilogue
        // adds the eight partial logits up
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int a = ((b & 1) << 2) | k;
          umma_bf16(tmem_base + (fs * kAcc + a) * kN, umma_smem_desc(xa + k * 32), umma_smem_desc(wa + k * 32),
                    idesc_fwd, b >= 2 ? 1u : 0u);
        }
        if (b == nb - 1) umma_commit(f_full + fs);
      }
      __syncwarp();
    };
    // The forward product runs ahead of the gradient product by what the ring holds beside one tile,
    // at most one whole tile (there are two forward accumulators): with that much lead the next tile's
    // coefficients are usually ready when this tile's gradient has been issued.
    const uint64_t total = static_cast<uint64_t>(my_tiles) * nb;
    const uint64_t lead = static_cast<uint64_t>(look < nb ? look : nb);
    uint64_t fseq = 0;
    int64_t ft = 0;   // tile and block of the next forward block
    int fb0 = 0;
    uint32_t gs0 = 0;  // ring slot of the first block of the tile whose gradient comes next
    for (int64_t i = 0; i < my_tiles; ++i) {
      uint64_t target = static_cast<uint64_t>(i + 1) * nb + lead;
      if (target > total) target = total;
      for (; fseq < target; ++fseq) {
        forward_block(ft, fb0);
        if (++fb0 == nb) { fb0 = 0; ++ft; }
      }
      // gradient of tile i: needs its coef operand
      const uint32_t cs = static_cast<uint32_t>(i & 1);
      mbar_wait(c_full + cs, (i >> 1) & 1);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t ca = smem_u32(sC + cs * kCoefBytes);
        // K-step outer, feature group inner: consecutive MMAs hit different accumulators
        for (int k = 0; k < kRows / 16; ++k) {
          for (int fb = 0; fb < nfb; ++fb) {
            uint32_t s = gs0 + 2 * fb;  // even, and s + 1 < R: the pair is adjacent
            if (s >= static_cast<uint32_t>(R)) s -= R;
            const uint32_t xa = smem_u32(sX + s * kBlk);
            umma_bf16(tmem_grad + fb * kN, umma_smem_desc_mn(xa + k * 2048, kBlk, 1024),
                      umma_smem_desc(ca + k * 32), idesc_grad, (i | k) ? 1u : 0u);
          }
        }
        for (int b = 0; b < nb; ++b) {
          uint32_t s = gs0 + b;
          if (s >= static_cast<uint32_t>(R)) s -= R;
          umma_commit(x_empty + s);
        }
        umma_commit(c_empty + cs);
        if (i == my_tiles - 1) umma_commit(g_full);
      }
      gs0 += static_cast<uint32_t>(nb);
      if (gs0 >= static_cast<uint32_t>(R)) gs0 -= R;
      __syncwarp();
    }
  } else {
    // ======================================================= epilogue
    const int q = warp & 3;  // TMEM lane quarter this warp may touch
    float loss = 0.f, correct = 0.f;
    if (q < kRows / 32) {
      for (int64_t i = 0; i < my_tiles; ++i) {
        const int64_t tile = blockIdx.x + i * gridDim.x;
        const uint32_t fs = static_cast<uint32_t>(i & 1);
        mbar_wait(f_full + fs, (i >> 1) & 1);
        tc_fence_after();
        float zsum = 0.f;
#pragma unroll
        for (int a = 0; a < kAcc; ++a) {
          uint32_t r1;
          asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];"
                       : "=r"(r1)
                       : "r"(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + (fs * kAcc + a) * kN)
                       : "memory");
          tmem_ld_wait();
          zsum += __uint_as_float(r1);  // column 0 of accumulator a
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(f_empty + fs);
        const int k = q * 32 + lane;  // sample inside the tile
        const int64_t row = tile * kRows + k;
        float coef = 0.f;
        if (row < p.n) {
          const float z = zsum;  // x . w
          const float y = p.label[row];
          const float pred = 1.f / (1.f + __expf(-z));
          coef = p.alpha * (y - pred);
          loss += fmaxf(z, 0.f) + log1pf(__expf(-fabsf(z))) - y * z;
          correct += ((pred > 0.5f) == (y > 0.5f)) ? 1.f : 0.f;
        }
        // coef as hi + lo bf16 into rows 0 / 1 of the K-major, 128B-swizzled B operand of the gradient
        mbar_wait(c_empty + fs, ((i >> 1) & 1) ^ 1);
        const __nv_bfloat16 hi = __float2bfloat16_rn(coef);
        const __nv_bfloat16 lo = __float2bfloat16_rn(coef - __bfloat162float(hi));
        uint8_t* cb = sC + fs * kCoefBytes;
        const int chunk = k >> 3, within = (k & 7) * 2;
        *reinterpret_cast<__nv_bfloat16*>(cb + 0 * 128 + ((chunk ^ 0) << 4) + within) = hi;
        *reinterpret_cast<__nv_bfloat16*>(cb + 1 * 128 + ((chunk ^ 1) << 4) + within) = lo;
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(c_full + fs);
      }
      loss = warp_sum(loss);
      correct = warp_sum(correct);
      if (lane == 0 && p.loss_out) {
        atomicAdd(p.loss_out, loss);
        atomicAdd(p.loss_out + 1, correct);
      }
    }
    // gradient of everything this CTA saw: feature fb*128 + lane of lane quarter q, columns 0 (hi) + 1 (lo)
    if (my_tiles > 0) {
      mbar_wait(g_full, 0);
      tc_fence_after();
      for (int fb = 0; fb < nfb; ++fb) {
        uint32_t r[16];
        tmem_ld16(tmem_grad + (static_cast<uint32_t>(q * 32) << 16) + fb * kN, r);
        tmem_ld_wait();
        const int d = fb * 128 + q * 32 + lane;
        const float g = __uint_as_float(r[0]) + __uint_as_float(r[1]);
        if (d < p.D && g != 0.f) red_gpu(p.grad + d, g);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, kTmemCols);
}

// 2-D bf16 row-major [rows, cols] tensor, box = 64 columns x box_rows rows, 128B swizzle
static int make_map(CUtensorMap* map, const void* base, int64_t rows, int64_t cols, int box_rows) {
  return make_map_bf16(map, base, rows, cols, box_rows);
}

}  // namespace dlrf
}  // namespace mps

extern "C" int mps_dense_lr_step_tc(const MpsTableView* weights_view, const MpsClockView* clk, int progress,
                                    const float* w_f32, const void* x_bf16, const float* label, int64_t n, int D,
                                    float alpha, void* w16_scratch, float* w_f32_out, int* sync_ctr, float* grad,
                                    float* loss_out, int max_ctas, void* stream) {
  if (n <= 0) return 0;
  using namespace mps::dlrf;  // NOLINT
  const int Dp = (D + 127) / 128 * 128;
  if (D <= 0 || D % 8 != 0 || Dp / 64 > kMaxBlocks) return static_cast<int>(cudaErrorInvalidValue);
  Params p{};
  if (weights_view) {
    p.weights = *weights_view;
    p.has_view = 1;
    if (weights_view->dtype != MPS_F32) return static_cast<int>(cudaErrorInvalidValue);
  }
  p.w_f32 = w_f32;
  p.gated = (clk != nullptr && weights_view != nullptr) ? 1 : 0;
  if (p.gated) p.clk = *clk;
  p.progress = progress;
  p.w16 = static_cast<__nv_bfloat16*>(w16_scratch);
  p.w_f32_out = w_f32_out;
  p.sync_ctr = sync_ctr;
  p.label = label;
  p.alpha = alpha;
  p.grad = grad;
  p.loss_out = loss_out;
  p.n = n;
  p.D = D;
  p.Dp = Dp;
  const int nb = Dp / 64;
  // shared memory: ring + one block of slack + w blocks + 2 coef operands + barriers
  const int fixed = kSlack * kBlk + nb * kWBlk + 2 * kCoefBytes + 8 * (2 * kMaxRing + 10) + 16 + 1024;
  int ring = (227 * 1024 - fixed) / kBlk;
  if (ring > kMaxRing) ring = kMaxRing;
  ring &= ~1;  // feature pairs of blocks must not wrap around the ring
  if (ring < nb + 2) return static_cast<int>(cudaErrorInvalidValue);
  p.ring = ring;
  CUtensorMap map_x, map_w;
  int rc = make_map(&map_x, x_bf16, n, D, kRows);  // rows / columns beyond the matrix are zero-filled by TMA
  if (rc) return rc;
  rc = make_map(&map_w, w16_scratch, kN, Dp, kN);
  if (rc) return rc;
  const int smem = ring * kBlk + fixed;
  MPS_CHECK(cudaFuncSetAttribute(dense_lr_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int64_t n_tiles = (n + kRows - 1) / kRows;
  int grid = static_cast<int>(n_tiles < sms ? n_tiles : sms);
  if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
  auto st = static_cast<cudaStream_t>(stream);
  MPS_CHECK(cudaMemsetAsync(sync_ctr, 0, sizeof(int), st));
  void* args[] = {&map_x, &map_w, &p};
  MPS_CHECK(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(dense_lr_fused_kernel), dim3(grid), dim3(kThreads),
                                        args, smem, st));
  return 0;
}

// this translation unit's mbarrier time-out flag (tc_common.cuh); mps_dense_lr_tc_error() collects them
extern "C" int mps_dense_lr_fused_timeout(void) { return mps::tc::tc_take_timeout_flag(); }
