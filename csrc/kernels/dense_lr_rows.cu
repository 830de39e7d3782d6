// dense_lr_rows.cu — one dense logistic-regression iteration that reads X once and keeps NO tile resident.
//
// dense_lr_stream.cu / dense_lr_fused.cu hold a 64-row tile in shared memory from its first block until
// its gradient is done, because their gradient phase maps threads to FEATURES and so needs every row's
// coefficient first: 128 KiB of the 227 KiB are pinned and the copy of the next tile cannot overlap the
// math of this one (0.64 ms at 1 Mi x 1024 where the copy alone takes 0.33 ms).  Here one WARP owns one
// ROW from the load to the gradient:
//   lane l holds the weights of 32 features (16-byte chunks l, l+32, l+64, l+96 of a row) and 32 gradient
//   accumulators in registers for the whole kernel; it reads its 4 chunks of the row once from shared
//   memory, the warp reduces the dot product with shuffles, every lane evaluates sigmoid / loss / coef, and
//   the gradient update uses the SAME registers the forward just read.
// Nothing is exchanged between warps per tile (no block barrier, no coefficient array), a row is dead as
// soon as its warp is done with it, and the ring is just bytes in flight: 30-row stages filled by ONE
// linear bulk copy each (rows of a row-major matrix are contiguous, so no tensor map and no swizzle: the 32
// lanes of a warp read 512 contiguous bytes per instruction), 3 stages at D = 1024.
//   warp 0      producer: cp.async.bulk global -> shared, completion on the stage's mbarrier;
//   warps 1-15  consumers: 2 rows per warp and stage (16 warps in all, so every thread has 128 registers
//               for its 32 weights, 32 accumulators and the row in flight).
// Measured at 1 Mi x 1024 bf16 on one B200 (tools/dense_lr_modes.sh, whole iteration incl. Get / Add /
// Clock): 0.349 ms = 6.15 TB/s = 0.94 of the measured HBM copy rate (stream 0.64 ms, tcgen05 0.66 ms, PyTorch
// 0.72 ms).  About 180 instructions per row and lane, half of the issue slots at that rate.
// At the end the 15 warps' accumulators meet in shared memory (the ring is free by then) and leave with one
// red.add per feature and CTA.
// Reference op: apps/logistic_regression/lr_example.cpp:293-312 for dense rows.
// Limits: D % 8 == 0, D <= 1024 (32 lanes x 4 chunks x 8 features).
#include "tc_common.cuh"

namespace mps {
namespace dlrr {

using namespace mps::tc;  // NOLINT

constexpr int kComputeWarps = 15;                     // + the producer = 16 warps: 128 registers per thread
constexpr int kThreads = 32 * (1 + kComputeWarps);
constexpr int kRPW = 2;                               // rows per warp and stage (the label hand-over assumes 2)
constexpr int kTileRows = kComputeWarps * kRPW;       // 30
static_assert(kRPW == 2, "the per-row label pick below assumes two rows per warp");
constexpr int kMaxStages = 8;
constexpr uint32_t kPiece = 16384;                    // bytes per bulk copy

struct Params {
  MpsTableView weights;   // fp32 table of D keys
  int has_view;
  const float* w_f32;     // local dense weights when has_view == 0
  MpsClockView clk;
  int gated, progress;
  float* w_f32_out;       // optional [D]
  const uint8_t* x;       // [n, D] bf16 row-major
  const float* label;     // [n] in {0,1}
  float alpha;
  float* grad;            // [D] += sum_r coef[r] x[r, :]
  float* loss_out;        // [2]: += logloss sum, += #correct
  int64_t n;
  int D;
  int stages;
  uint32_t stage_bytes;   // kTileRows * D * 2 rounded up to 128
};

__device__ __forceinline__ void bulk_load(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "r"(addr)
               : "memory");
  return v;
}

__global__ void __launch_bounds__(kThreads, 1) dense_lr_rows_kernel(const __grid_constant__ Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  const int S = p.stages;
  uint8_t* sX = smem;
  uint64_t* full = reinterpret_cast<uint64_t*>(sX + static_cast<size_t>(S) * p.stage_bytes);
  uint64_t* empty = full + kMaxStages;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (p.gated) cta_gate(p.clk, p.progress);
  if (threadIdx.x == 0) {
    for (int i = 0; i < S; ++i) { mbar_init(full + i, 1); mbar_init(empty + i, kComputeWarps); }
    fence_barrier_init();
  }
  __syncthreads();
  const int64_t n_tiles = (p.n + kTileRows - 1) / kTileRows;
  const int64_t my_tiles = n_tiles > blockIdx.x ? (n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const uint32_t row_bytes = static_cast<uint32_t>(p.D) * 2;

  if (warp == 0) {
    if (elect_one()) {
      uint32_t s = 0, ph = 0;
      for (int64_t i = 0; i < my_tiles; ++i) {
        const int64_t row0 = (blockIdx.x + i * gridDim.x) * kTileRows;
        const int rows = p.n - row0 < kTileRows ? static_cast<int>(p.n - row0) : kTileRows;
        const uint32_t bytes = static_cast<uint32_t>(rows) * row_bytes;
        mbar_wait(empty + s, ph ^ 1);
        mbar_expect_tx(full + s, bytes);
        const uint8_t* src = p.x + row0 * row_bytes;
        uint8_t* dst = sX + static_cast<size_t>(s) * p.stage_bytes;
        for (uint32_t off = 0; off < bytes; off += kPiece)
          bulk_load(dst + off, src + off, bytes - off < kPiece ? bytes - off : kPiece, full + s);
        if (++s == static_cast<uint32_t>(S)) { s = 0; ph ^= 1; }
      }
    }
    return;
  }

  // ---- consumer warps
  const int cw = warp - 1;  // 0..14
  const int nch = p.D / 8;  // 16-byte chunks per row
  // the Get, fused: every lane pulls the weights of its own 32 features straight from the owning shards
  float w[4][8], g[4][8];
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
      g[c][j] = 0.f;
    }
  }
  float loss = 0.f, correct = 0.f;
  // labels come one stage ahead through the read-only path (4 bytes per 2 KiB row)
  float y_next[kRPW];
  auto fetch_labels = [&](int64_t i) {
    const int64_t r0 = (blockIdx.x + i * gridDim.x) * kTileRows + cw * kRPW;
#pragma unroll
    for (int rr = 0; rr < kRPW; ++rr) y_next[rr] = (i < my_tiles && r0 + rr < p.n) ? __ldg(p.label + r0 + rr) : 0.f;
  };
  fetch_labels(0);

  // everything that does not change per row is computed once: which of my 4 chunks exist, and the byte
  // offset of my first row's first chunk inside a stage (32-bit shared address arithmetic from here on)
  bool on[4];
#pragma unroll
  for (int c = 0; c < 4; ++c) on[c] = lane + 32 * c < nch;
  const uint32_t sx = smem_u32(sX);
  const uint32_t my_off = static_cast<uint32_t>(cw * kRPW) * row_bytes + static_cast<uint32_t>(lane) * 16;
  const float alpha = p.alpha;

  uint32_t s = 0, ph = 0;
  for (int64_t i = 0; i < my_tiles; ++i) {
    const int64_t left = p.n - (blockIdx.x + i * gridDim.x) * kTileRows;
    const int rows_here = left < kTileRows ? static_cast<int>(left) : kTileRows;
    float y[kRPW];
#pragma unroll
    for (int rr = 0; rr < kRPW; ++rr) y[rr] = y_next[rr];
    fetch_labels(i + 1);
    mbar_wait(full + s, ph);
    uint32_t addr = sx + s * p.stage_bytes + my_off;
#pragma unroll 1
    for (int rr = 0; rr < kRPW; ++rr, addr += row_bytes) {
      const bool live = cw * kRPW + rr < rows_here;  // warp-uniform; false only in the last tile of the batch
      uint4 x[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        x[c] = make_uint4(0u, 0u, 0u, 0u);
        if (on[c] && live) x[c] = lds128(addr + 512 * c);
      }
      float acc[4] = {0.f, 0.f, 0.f, 0.f};  // one chain per chunk
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const uint4 v = x[c];
        acc[c] = fmaf(__uint_as_float(v.x << 16), w[c][0], acc[c]);
        acc[c] = fmaf(__uint_as_float(v.x & 0xffff0000u), w[c][1], acc[c]);
        acc[c] = fmaf(__uint_as_float(v.y << 16), w[c][2], acc[c]);
        acc[c] = fmaf(__uint_as_float(v.y & 0xffff0000u), w[c][3], acc[c]);
        acc[c] = fmaf(__uint_as_float(v.z << 16), w[c][4], acc[c]);
        acc[c] = fmaf(__uint_as_float(v.z & 0xffff0000u), w[c][5], acc[c]);
        acc[c] = fmaf(__uint_as_float(v.w << 16), w[c][6], acc[c]);
        acc[c] = fmaf(__uint_as_float(v.w & 0xffff0000u), w[c][7], acc[c]);
      }
      const float z = warp_sum((acc[0] + acc[1]) + (acc[2] + acc[3]));
      const float yr = rr == 0 ? y[0] : y[kRPW - 1];
      // one exponential serves the sigmoid and the loss: e = exp(-|z|), sigmoid = (z >= 0 ? 1 : e) / (1 + e),
      // logloss = max(z, 0) - y z + log(1 + e)
      const float e = __expf(-fabsf(z));
      const float inv = __fdividef(1.f, 1.f + e);
      const float pred = z >= 0.f ? inv : e * inv;
      float coef = 0.f;
      if (live) {
        coef = alpha * (yr - pred);
        loss += fmaxf(z, 0.f) - yr * z + __logf(1.f + e);
        correct += ((pred > 0.5f) == (yr > 0.5f)) ? 1.f : 0.f;
      }
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const uint4 v = x[c];
        g[c][0] = fmaf(coef, __uint_as_float(v.x << 16), g[c][0]);
        g[c][1] = fmaf(coef, __uint_as_float(v.x & 0xffff0000u), g[c][1]);
        g[c][2] = fmaf(coef, __uint_as_float(v.y << 16), g[c][2]);
        g[c][3] = fmaf(coef, __uint_as_float(v.y & 0xffff0000u), g[c][3]);
        g[c][4] = fmaf(coef, __uint_as_float(v.z << 16), g[c][4]);
        g[c][5] = fmaf(coef, __uint_as_float(v.z & 0xffff0000u), g[c][5]);
        g[c][6] = fmaf(coef, __uint_as_float(v.w << 16), g[c][6]);
        g[c][7] = fmaf(coef, __uint_as_float(v.w & 0xffff0000u), g[c][7]);
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(empty + s);
    if (++s == static_cast<uint32_t>(S)) { s = 0; ph ^= 1; }
  }
  if (lane == 0 && p.loss_out && (loss != 0.f || correct != 0.f)) {
    atomicAdd(p.loss_out, loss);
    atomicAdd(p.loss_out + 1, correct);
  }
  // every stage this CTA asked for has landed and been read: the ring is scratch now.  [15 warps][D] partial
  // gradients, then one red.add per feature
  asm volatile("bar.sync 1, %0;" ::"n"(32 * kComputeWarps) : "memory");
  float* part = reinterpret_cast<float*>(sX);
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    if (lane + 32 * c < nch) {
      float4* dst = reinterpret_cast<float4*>(part + static_cast<size_t>(cw) * p.D + (lane + 32 * c) * 8);
      dst[0] = make_float4(g[c][0], g[c][1], g[c][2], g[c][3]);
      dst[1] = make_float4(g[c][4], g[c][5], g[c][6], g[c][7]);
    }
  }
  asm volatile("bar.sync 1, %0;" ::"n"(32 * kComputeWarps) : "memory");
  for (int d = cw * 32 + lane; d < p.D; d += 32 * kComputeWarps) {
    float sum = 0.f;
#pragma unroll
    for (int k = 0; k < kComputeWarps; ++k) sum += part[k * p.D + d];
    if (sum != 0.f) red_gpu(p.grad + d, sum);
  }
}

}  // namespace dlrr
}  // namespace mps

extern "C" int mps_dense_lr_step_rows(const MpsTableView* weights_view, const MpsClockView* clk, int progress,
                                      const float* w_f32, const void* x_bf16, const float* label, int64_t n, int D,
                                      float alpha, float* w_f32_out, float* grad, float* loss_out, int max_ctas,
                                      void* stream) {
  if (n <= 0) return 0;
  using namespace mps::dlrr;  // NOLINT
  if (D <= 0 || D % 8 != 0 || D > 1024) return static_cast<int>(cudaErrorInvalidValue);
  if (reinterpret_cast<uintptr_t>(x_bf16) & 15) return static_cast<int>(cudaErrorInvalidValue);
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
  p.x = static_cast<const uint8_t*>(x_bf16);
  p.label = label;
  p.alpha = alpha;
  p.grad = grad;
  p.loss_out = loss_out;
  p.n = n;
  p.D = D;
  p.stage_bytes = (static_cast<uint32_t>(kTileRows) * D * 2 + 127) & ~127u;
  const int fixed = 2 * 8 * kMaxStages + 128;
  int stages = (216 * 1024 - fixed) / static_cast<int>(p.stage_bytes);
  if (stages > kMaxStages) stages = kMaxStages;
  if (stages < 2) return static_cast<int>(cudaErrorInvalidValue);
  p.stages = stages;
  const int smem = stages * static_cast<int>(p.stage_bytes) + fixed;
  MPS_CHECK(cudaFuncSetAttribute(dense_lr_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int64_t n_tiles = (n + kTileRows - 1) / kTileRows;
  int grid = static_cast<int>(n_tiles < sms ? n_tiles : sms);
  if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
  dense_lr_rows_kernel<<<grid, kThreads, smem, static_cast<cudaStream_t>(stream)>>>(p);
  MPS_LAUNCH_CHECK();
  return 0;
}

// this translation unit's mbarrier time-out flag (tc_common.cuh); mps_dense_lr_tc_error() collects them
extern "C" int mps_dense_lr_rows_timeout(void) { return mps::tc::tc_take_timeout_flag(); }
