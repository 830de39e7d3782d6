// dense_lr_tc.cu — dense logistic-regression batches on the tensor cores.
//
// The reference's LR (apps/logistic_regression/lr_example.cpp:293-312) walks sparse samples; for
// DENSE feature batches the same iteration is a matrix product.  Forward: logits = X . w with X
// [n, D] bf16 streamed by TMA in 128-row x 64-column swizzled tiles, w held as row 0 of a 16-row
// B operand (the smallest UMMA N), tcgen05.mma accumulating over D into TMEM, and the
// sigmoid / logloss / alpha (y - p) epilogue applied straight out of TMEM — the logits never
// touch memory.  The weight PULL is fused into the prologue (every CTA reads its slice of the
// fp32 table from the owning shards and rounds it into the B operand).  Backward
// (grad = coef^T X) is a bandwidth-bound column reduction on the CUDA cores feeding the dense
// push.
#include "tc_common.cuh"

namespace mps {
namespace dlr {

using namespace mps::tc;  // NOLINT

constexpr int kTileM = 128;
constexpr int kN = 16;         // UMMA N (row 0 = w, rows 1..15 zero)
constexpr int kStages = 4;
constexpr int kSlots = 2;      // TMEM accumulator slots of kN columns
constexpr int kThreads = 192;  // warp 0 TMA, warp 1 MMA, warps 2..5 epilogue
constexpr int kABytes = kTileM * 128;  // [128 x 64] bf16
constexpr int kBBytes = kN * 128;      // [16 x 64] bf16
constexpr int kStageBytes = kABytes + kBBytes;

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}

struct Params {
  MpsTableView weights;   // fp32 table of D keys
  int has_view;
  const float* w_f32;     // local dense weights when has_view == 0
  MpsClockView clk;
  int gated, progress;
  __nv_bfloat16* w16;     // scratch [16, Dp] bf16 (Dp = D rounded up to 64)
  float* w_f32_out;       // optional [D]
  int* sync_ctr;
  const float* label;     // [n] in {0,1}
  float alpha;
  float* coef;            // [n] out: alpha * (y - sigmoid(x.w))
  float* loss_out;        // [2]: += logloss sum, += #correct
  int64_t n;
  int D, Dp;
};

__global__ void __launch_bounds__(kThreads, 1)
dense_lr_forward_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w,
                        const __grid_constant__ Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * kStageBytes);
  uint64_t* s_full = bars;
  uint64_t* s_empty = s_full + kStages;
  uint64_t* t_full = s_empty + kStages;
  uint64_t* t_empty = t_full + kSlots;
  uint32_t* tmem_base_smem = reinterpret_cast<uint32_t*>(t_empty + kSlots);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // ---- prologue: pull w, round to bf16 into row 0 of the B operand (rows 1..15 and the pad stay zero)
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

  if (threadIdx.x == 0) {
    for (int i = 0; i < kStages; ++i) { mbar_init(s_full + i, 1); mbar_init(s_empty + i, 1); }
    for (int i = 0; i < kSlots; ++i) { mbar_init(t_full + i, 1); mbar_init(t_empty + i, 4); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_base_smem, 32);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_base_smem;
  const int64_t n_tiles = (p.n + kTileM - 1) / kTileM;
  const int n_kc = p.Dp / 64;

  if (warp == 0) {
    if (elect_one()) {
      uint32_t si = 0;
      for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        for (int kc = 0; kc < n_kc; ++kc, ++si) {
          const uint32_t st = si % kStages;
          uint8_t* stage = smem + st * kStageBytes;
          mbar_wait(s_empty + st, ((si / kStages) & 1) ^ 1);
          mbar_expect_tx(s_full + st, kStageBytes);
          tma_load_2d(stage, &map_x, s_full + st, kc * 64, static_cast<int>(tile * kTileM));
          tma_load_2d(stage + kABytes, &map_w, s_full + st, kc * 64, 0);
        }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = umma_idesc(kTileM, kN);
    uint32_t si = 0, ti = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++ti) {
      const uint32_t slot = ti % kSlots;
      mbar_wait(t_empty + slot, ((ti / kSlots) & 1) ^ 1);
      for (int kc = 0; kc < n_kc; ++kc, ++si) {
        const uint32_t st = si % kStages;
        const uint32_t stage = smem_u32(smem + st * kStageBytes);
        mbar_wait(s_full + st, (si / kStages) & 1);
        tc_fence_after();
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_bf16(tmem_base + slot * kN, umma_smem_desc(stage + k * 32), umma_smem_desc(stage + kABytes + k * 32),
                      idesc, (kc | k) ? 1u : 0u);
          umma_commit(s_empty + st);
          if (kc == n_kc - 1) umma_commit(t_full + slot);
        }
        __syncwarp();
      }
    }
  } else {
    const int quarter = warp & 3;
    uint32_t ti = 0;
    float loss = 0.f, correct = 0.f;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++ti) {
      const uint32_t slot = ti % kSlots;
      mbar_wait(t_full + slot, (ti / kSlots) & 1);
      tc_fence_after();
      uint32_t r[16];
      tmem_ld16(tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + slot * kN, r);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(t_empty + slot);
      const int64_t row = tile * kTileM + quarter * 32 + lane;
      if (row < p.n) {
        const float z = __uint_as_float(r[0]);  // column 0 = x . w
        const float y = p.label[row];
        const float pred = 1.f / (1.f + __expf(-z));
        p.coef[row] = p.alpha * (y - pred);
        loss += fmaxf(z, 0.f) + log1pf(__expf(-fabsf(z))) - y * z;
        correct += ((pred > 0.5f) == (y > 0.5f)) ? 1.f : 0.f;
      }
    }
    loss = warp_sum(loss);
    correct = warp_sum(correct);
    if (lane == 0 && p.loss_out) {
      atomicAdd(p.loss_out, loss);
      atomicAdd(p.loss_out + 1, correct);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 32);
}

// grad[d] += sum_r coef[r] * x[r, d]: CTA = 128 rows x 512 columns, 2 columns per thread
__global__ void __launch_bounds__(256)
dense_lr_grad_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ coef, int64_t n, int D,
                     float* __restrict__ grad) {
  __shared__ float s_coef[128];
  const int64_t row0 = static_cast<int64_t>(blockIdx.y) * 128;
  const int col = (blockIdx.x * 256 + threadIdx.x) * 2;
  for (int i = threadIdx.x; i < 128; i += 256) s_coef[i] = (row0 + i < n) ? coef[row0 + i] : 0.f;
  __syncthreads();
  if (col >= D) return;
  float a0 = 0.f, a1 = 0.f;
  const int rows = static_cast<int>(n - row0 < 128 ? n - row0 : 128);
  if (col + 1 < D && (D & 1) == 0) {
#pragma unroll 4
    for (int r = 0; r < rows; ++r) {
      const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(x + (row0 + r) * D + col);
      const float2 f = __bfloat1622float2(v);
      a0 = fmaf(s_coef[r], f.x, a0);
      a1 = fmaf(s_coef[r], f.y, a1);
    }
    if (a0 != 0.f) red_gpu(grad + col, a0);
    if (a1 != 0.f) red_gpu(grad + col + 1, a1);
  } else {
    for (int c = col; c < D && c < col + 2; ++c) {
      float a = 0.f;
      for (int r = 0; r < rows; ++r) a = fmaf(s_coef[r], __bfloat162float(x[(row0 + r) * D + c]), a);
      if (a != 0.f) red_gpu(grad + c, a);
    }
  }
}

}  // namespace dlr
}  // namespace mps

extern "C" {

int mps_dense_lr_forward_tc(const MpsTableView* weights_view, const MpsClockView* clk, int progress,
                            const float* w_f32, const void* x_bf16, const float* label, int64_t n, int D, float alpha,
                            void* w16_scratch, float* w_f32_out, int* sync_ctr, float* coef, float* loss_out,
                            int max_ctas, void* stream) {
  if (n <= 0) return 0;
  if (D <= 0 || D % 8 != 0) return static_cast<int>(cudaErrorInvalidValue);  // TMA row pitch must be 16-byte aligned
  using namespace mps::dlr;  // NOLINT
  Params p{};
  if (weights_view) {
    p.weights = *weights_view;
    p.has_view = 1;
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
  p.coef = coef;
  p.loss_out = loss_out;
  p.n = n;
  p.D = D;
  p.Dp = (D + 63) / 64 * 64;
  CUtensorMap map_x, map_w;
  int rc = mps::tc::make_map_bf16(&map_x, x_bf16, n, D, kTileM);  // columns beyond D are zero-filled by TMA
  if (rc) return rc;
  rc = mps::tc::make_map_bf16(&map_w, w16_scratch, kN, p.Dp, kN);
  if (rc) return rc;
  const int smem = kStages * kStageBytes + 8 * (2 * kStages + 2 * kSlots) + 16 + 1024;
  MPS_CHECK(cudaFuncSetAttribute(dense_lr_forward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int64_t n_tiles = (n + kTileM - 1) / kTileM;
  int grid = static_cast<int>(n_tiles < sms ? n_tiles : sms);
  if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
  auto st = static_cast<cudaStream_t>(stream);
  MPS_CHECK(cudaMemsetAsync(sync_ctr, 0, sizeof(int), st));
  void* args[] = {&map_x, &map_w, &p};
  MPS_CHECK(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(dense_lr_forward_kernel), dim3(grid), dim3(kThreads),
                                        args, smem, st));
  return 0;
}

int mps_dense_lr_grad(const void* x_bf16, const float* coef, int64_t n, int D, float* grad, void* stream) {
  if (n <= 0) return 0;
  dim3 grid(static_cast<unsigned>((D + 511) / 512), static_cast<unsigned>((n + 127) / 128));
  mps::dlr::dense_lr_grad_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x_bf16), coef, n, D, grad);
  MPS_LAUNCH_CHECK();
  return 0;
}

/* 1 if an mbarrier wait of a tcgen05 kernel of this file expired since the last call (clears the flag);
 * synchronises with the device. */
int mps_dense_lr_fused_timeout(void);
int mps_dense_lr_stream_timeout(void);
int mps_dense_lr_rows_timeout(void);
// every dense-LR translation unit has its own copy of the flag
int mps_dense_lr_tc_error(void) {
  const int a = mps::tc::tc_take_timeout_flag(), b = mps_dense_lr_fused_timeout(), c = mps_dense_lr_stream_timeout(),
            d = mps_dense_lr_rows_timeout();
  return (a < 0 || b < 0 || c < 0 || d < 0) ? -1 : (a | b | c | d);
}

}  // extern "C"
