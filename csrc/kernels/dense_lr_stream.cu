// dense_lr_stream.cu — one dense logistic-regression iteration that reads X once, at the HBM rate.
//
// A dense LR step is two matrix-VECTOR products over the same batch (logits = X w, grad = coef^T X): two
// flops per byte of X each, so the only thing that matters is how often X crosses HBM.  dense_lr_tc.cu
// reads it twice (tcgen05 forward, then a CUDA-core gradient pass).  Two kernels read it once:
// dense_lr_fused.cu keeps both products on the tensor cores (the resident tile doubles as an MN-major A
// operand, the gradient accumulates in TMEM); this one feeds the CUDA cores from the same TMA ring, which
// a matrix-vector product (no operand reuse) can afford.  Measured at 1 Mi x 1024 bf16 on one B200
// (tools/dense_lr_modes.sh, whole iteration incl. Get / Add / Clock): this kernel 0.64 ms, the tcgen05
// one 0.66 ms, two passes 0.75 ms, PyTorch (two cuBLAS GEMVs) 0.73 ms.  Both one-pass kernels are held
// back by residency, not by HBM (0.33 ms at the copy rate): a 64-row tile must stay in shared memory from
// its first block until its gradient is done, and 227 KiB hold one tile plus 6-10 blocks of the next.
// dense_lr_rows.cu removes the residency (one warp owns a row end to end) and runs at 0.35 ms; it is the
// default, this kernel and the tcgen05 one stay selectable (MINIPS_DENSE_LR).
//   warp 0      TMA producer: 64-row x 64-column blocks of X (bf16, 128B swizzle) into a ring that holds
//               one tile plus the first blocks of the next;
//   warps 1-16  per tile: forward  - one warp per row, every lane owns 32 features whose weights stay in
//                                    registers for the whole kernel; sigmoid / logloss / coef on the spot;
//                         gradient - every thread owns 4 features and walks 32 of the 64 rows of the SAME
//                                    shared-memory tile (two threads per feature group); its accumulators
//                                    stay in registers for the whole kernel and leave with one red.add
//                                    per feature and thread.
// All shared-memory reads follow the 128B swizzle of the TMA blocks and are bank-conflict free.
// Reference op: apps/logistic_regression/lr_example.cpp:293-312 for dense rows.
// Limits: D % 8 == 0 and roundup(D, 128) <= 1024 (thread / register mapping: 512 threads x 4 features, 32
// lanes x 32 weights).
#include "tc_common.cuh"

namespace mps {
namespace dlrs {

using namespace mps::tc;  // NOLINT

constexpr int kRows = 64;           // rows of X per tile (32-row tiles, three of them resident, measured slower:
                                    // 0.96 ms against 0.64 ms per iteration at 1 Mi x 1024)
constexpr int kBlk = kRows * 128;   // [64 rows x 64 cols] bf16 = 8 KiB
constexpr int kComputeWarps = 16;
constexpr int kThreads = 32 * (1 + kComputeWarps);
constexpr int kMaxRing = 26;

struct Params {
  MpsTableView weights;   // fp32 table of D keys
  int has_view;
  const float* w_f32;     // local dense weights when has_view == 0
  MpsClockView clk;
  int gated, progress;
  float* w_f32_out;       // optional [D]
  const float* label;     // [n] in {0,1}
  float alpha;
  float* grad;            // [D] += sum_r coef[r] x[r, :]
  float* loss_out;        // [2]: += logloss sum, += #correct
  int64_t n;
  int D, Dp;              // Dp = D rounded up to 128
  int ring;
};

// 16-byte chunk `chunk` (8 features) of row r of a tile whose blocks live in ring slots s0, s0+1, ...
__device__ __forceinline__ uint32_t tile_chunk_off(int r, int chunk) {
  return static_cast<uint32_t>((chunk >> 3) * kBlk + r * 128 + (((chunk & 7) ^ (r & 7)) << 4));
}

__global__ void __launch_bounds__(kThreads, 1)
dense_lr_stream_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int R = p.ring;
  const int nb = p.Dp / 64;
  uint8_t* sX = smem;
  float* s_coef = reinterpret_cast<float*>(sX + static_cast<size_t>(R) * kBlk);  // [2][64]
  uint64_t* x_full = reinterpret_cast<uint64_t*>(s_coef + 2 * kRows);
  uint64_t* x_empty = x_full + kMaxRing;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (p.gated) cta_gate(p.clk, p.progress);
  if (threadIdx.x == 0) {
    for (int i = 0; i < R; ++i) { mbar_init(x_full + i, 1); mbar_init(x_empty + i, kComputeWarps); }
    fence_barrier_init();
  }
  __syncthreads();
  const int64_t n_tiles = (p.n + kRows - 1) / kRows;
  const int64_t my_tiles = n_tiles > blockIdx.x ? (n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;

  if (warp == 0) {
    if (elect_one()) {
      // ring position kept incrementally (a 64-bit divide per block would cost more than the block's math)
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
    return;
  }

  // ---- compute warps
  const int cw = warp - 1;              // 0..15
  const int ct = cw * 32 + lane;        // 0..511
  // the Get, fused: every lane pulls the weights of its own 32 features (4 chunks: lane, lane+32, ...)
  // straight from the owning shards; they stay in registers
  float w[4][8];
#pragma unroll
  for (int c = 0; c < 4; ++c) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int d = (lane + 32 * c) * 8 + j;
      float v = 0.f;
      if (d < p.D) {
        if (p.has_view) {
          const uint32_t key = p.weights.begin[0] + static_cast<uint32_t>(d);
          const int s = owner_of(p.weights, key);
          v = ld_sys(static_cast<const float*>(p.weights.weights[s]) + (key - p.weights.begin[s]));
        } else {
          v = p.w_f32[d];
        }
        if (p.w_f32_out && blockIdx.x == 0 && cw == 0) p.w_f32_out[d] = v;
        v = __bfloat162float(__float2bfloat16_rn(v));  // the same rounding as the tensor-core path
      }
      w[c][j] = v;
    }
  }
  // gradient of features 4*cf .. 4*cf+3 over one half of the rows of every tile of this CTA
  const int cf = ct & 255, rh = ct >> 8;
  float ga[4] = {0.f, 0.f, 0.f, 0.f}, gb[4] = {0.f, 0.f, 0.f, 0.f};
  float loss = 0.f, correct = 0.f;
  const int nchunks = p.Dp / 8;         // 16-byte chunks per row
  constexpr int kRowsPerWarp = kRows / kComputeWarps;  // 4

  uint32_t s0 = 0, ph0 = 0;  // ring slot and phase of the tile's first block
  auto slot_of = [&](int b) -> uint32_t {  // slot of block b of the current tile (b < nb <= R)
    const uint32_t t = s0 + static_cast<uint32_t>(b);
    return t >= static_cast<uint32_t>(R) ? t - R : t;
  };
  for (int64_t i = 0; i < my_tiles; ++i) {
    const int64_t tile = blockIdx.x + i * gridDim.x;
    if (lane == 0)
      for (int b = 0; b < nb; ++b) {
        const uint32_t t = s0 + static_cast<uint32_t>(b);
        const bool wrapped = t >= static_cast<uint32_t>(R);
        mbar_wait(x_full + (wrapped ? t - R : t), wrapped ? ph0 ^ 1 : ph0);
      }
    __syncwarp();
    float* coefs = s_coef + (i & 1) * kRows;
    // forward: rows cw*4 .. cw*4+3; lane l reads chunks l, l+32, l+64, l+96 of a row (one block each)
    const uint8_t* fb[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) fb[c] = sX + slot_of(((lane + 32 * c) >> 3) < nb ? ((lane + 32 * c) >> 3) : 0) * kBlk;
    float z[kRowsPerWarp];
#pragma unroll
    for (int rr = 0; rr < kRowsPerWarp; ++rr) {
      const int r = cw * kRowsPerWarp + rr;
      float acc[4] = {0.f, 0.f, 0.f, 0.f};  // one chain per chunk
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int chunk = lane + 32 * c;
        if (chunk < nchunks) {
          const uint4 v = *reinterpret_cast<const uint4*>(fb[c] + r * 128 + (((chunk & 7) ^ (r & 7)) << 4));
          acc[c] = fmaf(__uint_as_float(v.x << 16), w[c][0], acc[c]);
          acc[c] = fmaf(__uint_as_float(v.x & 0xffff0000u), w[c][1], acc[c]);
          acc[c] = fmaf(__uint_as_float(v.y << 16), w[c][2], acc[c]);
          acc[c] = fmaf(__uint_as_float(v.y & 0xffff0000u), w[c][3], acc[c]);
          acc[c] = fmaf(__uint_as_float(v.z << 16), w[c][4], acc[c]);
          acc[c] = fmaf(__uint_as_float(v.z & 0xffff0000u), w[c][5], acc[c]);
          acc[c] = fmaf(__uint_as_float(v.w << 16), w[c][6], acc[c]);
          acc[c] = fmaf(__uint_as_float(v.w & 0xffff0000u), w[c][7], acc[c]);
        }
      }
      z[rr] = warp_sum((acc[0] + acc[1]) + (acc[2] + acc[3]));
    }
    // lane rr finishes row rr of this warp: sigmoid, loss, coefficient
    if (lane < kRowsPerWarp) {
      float zz = z[0];
#pragma unroll
      for (int rr = 1; rr < kRowsPerWarp; ++rr) zz = lane == rr ? z[rr] : zz;
      const int r = cw * kRowsPerWarp + lane;
      const int64_t row = tile * kRows + r;
      float coef = 0.f;
      if (row < p.n) {
        const float y = p.label[row];
        const float pred = 1.f / (1.f + __expf(-zz));
        coef = p.alpha * (y - pred);
        loss += fmaxf(zz, 0.f) + log1pf(__expf(-fabsf(zz))) - y * zz;
        correct += ((pred > 0.5f) == (y > 0.5f)) ? 1.f : 0.f;
      }
      coefs[r] = coef;
    }
    // every row's coefficient is needed by every thread: barrier among the compute warps only
    asm volatile("bar.sync 1, %0;" ::"n"(32 * kComputeWarps) : "memory");
    // gradient: features 4*cf .. 4*cf+3 = half of chunk cf/2, rows of my half, two rows per step
    if (4 * cf < p.Dp) {
      const int chunk = cf >> 1, half = (cf & 1) * 8;
      const uint8_t* blk = sX + slot_of(chunk >> 3) * kBlk + half;
      const int c7 = chunk & 7;
#pragma unroll 4
      for (int r = rh * (kRows / 2); r < (rh + 1) * (kRows / 2); r += 2) {
        const uint2 va = *reinterpret_cast<const uint2*>(blk + r * 128 + ((c7 ^ (r & 7)) << 4));
        const uint2 vb = *reinterpret_cast<const uint2*>(blk + (r + 1) * 128 + ((c7 ^ ((r + 1) & 7)) << 4));
        const float ca = coefs[r], cb = coefs[r + 1];
        ga[0] = fmaf(ca, __uint_as_float(va.x << 16), ga[0]);
        ga[1] = fmaf(ca, __uint_as_float(va.x & 0xffff0000u), ga[1]);
        ga[2] = fmaf(ca, __uint_as_float(va.y << 16), ga[2]);
        ga[3] = fmaf(ca, __uint_as_float(va.y & 0xffff0000u), ga[3]);
        gb[0] = fmaf(cb, __uint_as_float(vb.x << 16), gb[0]);
        gb[1] = fmaf(cb, __uint_as_float(vb.x & 0xffff0000u), gb[1]);
        gb[2] = fmaf(cb, __uint_as_float(vb.y << 16), gb[2]);
        gb[3] = fmaf(cb, __uint_as_float(vb.y & 0xffff0000u), gb[3]);
      }
    }
    __syncwarp();
    if (lane == 0)
      for (int b = 0; b < nb; ++b) mbar_arrive(x_empty + slot_of(b));
    s0 += static_cast<uint32_t>(nb);
    if (s0 >= static_cast<uint32_t>(R)) { s0 -= R; ph0 ^= 1; }
  }
  loss = warp_sum(loss);
  correct = warp_sum(correct);
  if (lane == 0 && p.loss_out && (loss != 0.f || correct != 0.f)) {
    atomicAdd(p.loss_out, loss);
    atomicAdd(p.loss_out + 1, correct);
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int d = 4 * cf + j;
    const float gsum = ga[j] + gb[j];
    if (d < p.D && gsum != 0.f) red_gpu(p.grad + d, gsum);
  }
}

}  // namespace dlrs
}  // namespace mps

extern "C" int mps_dense_lr_step_stream(const MpsTableView* weights_view, const MpsClockView* clk, int progress,
                                        const float* w_f32, const void* x_bf16, const float* label, int64_t n, int D,
                                        float alpha, float* w_f32_out, float* grad, float* loss_out, int max_ctas,
                                        void* stream) {
  if (n <= 0) return 0;
  using namespace mps::dlrs;  // NOLINT
  const int Dp = (D + 127) / 128 * 128;
  if (D <= 0 || D % 8 != 0 || Dp > 1024) return static_cast<int>(cudaErrorInvalidValue);
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
  p.w_f32_out = w_f32_out;
  p.label = label;
  p.alpha = alpha;
  p.grad = grad;
  p.loss_out = loss_out;
  p.n = n;
  p.D = D;
  p.Dp = Dp;
  const int nb = Dp / 64;
  const int fixed = 2 * kRows * 4 + 8 * 2 * kMaxRing + 1024;
  int ring = (220 * 1024 - fixed) / kBlk;
  if (ring > kMaxRing) ring = kMaxRing;
  if (ring < nb + 1) return static_cast<int>(cudaErrorInvalidValue);
  p.ring = ring;
  CUtensorMap map_x;
  int rc = mps::tc::make_map_bf16(&map_x, x_bf16, n, D, kRows);  // rows / columns beyond the matrix read as zero
  if (rc) return rc;
  const int smem = ring * kBlk + fixed;
  MPS_CHECK(cudaFuncSetAttribute(dense_lr_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int64_t n_tiles = (n + kRows - 1) / kRows;
  int grid = static_cast<int>(n_tiles < sms ? n_tiles : sms);
  if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
  dense_lr_stream_kernel<<<grid, kThreads, smem, static_cast<cudaStream_t>(stream)>>>(map_x, p);
  MPS_LAUNCH_CHECK();
  return 0;
}

// this translation unit's mbarrier time-out flag (tc_common.cuh); mps_dense_lr_tc_error() collects them
extern "C" int mps_dense_lr_stream_timeout(void) { return mps::tc::tc_take_timeout_flag(); }
