// push_pull.cu — KVClientTable::Get / Add as one-sided device operations.
//
// Reference path (worker/kv_client_table.hpp:168-195): slice sorted keys per server, send
// kGet/kAdd over ZMQ, a server thread gathers / accumulates, the reply is sorted and
// memcpy-concatenated.  Here every key resolves its owning shard in-kernel and the value
// moves with a single ld / red on the owner's memory (local HBM or NVLink peer mapping);
// the "server-side +=" (server/vector_storage.hpp:28-38) is the red.add executed at the
// owner's L2.  Dense variants move whole shard slabs with 16-byte accesses.
#include "common.cuh"

namespace mps {

template <typename T>
__global__ void __launch_bounds__(256)
pull_kernel(const __grid_constant__ MpsTableView v, const __grid_constant__ MpsClockView c,
            int gated, int progress, const uint32_t* __restrict__ keys, T* __restrict__ vals,
            int64_t n) {
  if (gated) cta_gate(c, progress);
  const int k = v.vals_per_key;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint32_t key = keys[i];
    const int s = owner_of(v, key);
    if (s < 0) {
      for (int j = 0; j < k; ++j) vals[i * k + j] = T(0);
      continue;
    }
    const T* src = static_cast<const T*>(v.weights[s]) + static_cast<int64_t>(key - v.begin[s]) * k;
    for (int j = 0; j < k; ++j) vals[i * k + j] = ld_sys(src + j);
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
push_kernel(const __grid_constant__ MpsTableView v, const uint32_t* __restrict__ keys,
            const T* __restrict__ vals, int64_t n) {
  const int k = v.vals_per_key;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint32_t key = keys[i];
    const int s = owner_of(v, key);
    if (s < 0) continue;
    void* base = v.staging[s] ? v.staging[s] : v.weights[s];
    T* dst = static_cast<T*>(base) + static_cast<int64_t>(key - v.begin[s]) * k;
    for (int j = 0; j < k; ++j) {
      const T x = vals[i * k + j];
      if (x != T(0)) red_sys(dst + j, x);
    }
  }
}

// Flat element space of the table: shard s covers [fb(s), fb(s+1)) with
// fb(s) = (begin[s] - begin[0]) * k.  Work is rotated so that rank r starts on its own
// shard and all ranks hit different owners at any moment.
__device__ __forceinline__ int64_t flat_begin(const MpsTableView& v, int s) {
  return static_cast<int64_t>(v.begin[s] - v.begin[0]) * v.vals_per_key;
}

__global__ void __launch_bounds__(256)
pull_dense_f32_kernel(const __grid_constant__ MpsTableView v, const __grid_constant__ MpsClockView c,
                      int gated, int progress, int rot_shard, float* __restrict__ dst) {
  if (gated) cta_gate(c, progress);
  const int S = v.num_shards;
  const int64_t total = flat_begin(v, S);
  const int64_t rot = flat_begin(v, rot_shard) & ~int64_t(3);
  const int64_t nchunk = (total + 3) >> 2;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  const int k = v.vals_per_key;
  for (int64_t ci = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; ci < nchunk; ci += stride) {
    int64_t f = (ci << 2) + rot;
    if (f >= (nchunk << 2)) f -= (nchunk << 2);
    if (f >= total) continue;
    const uint32_t key0 = v.begin[0] + static_cast<uint32_t>(f / k);
    const int s = owner_of(v, key0);
    const int64_t fb = flat_begin(v, s), fe = flat_begin(v, s + 1);
    const float* src = static_cast<const float*>(v.weights[s]);
    if (f + 4 <= fe && ((f - fb) & 3) == 0) {
      *reinterpret_cast<float4*>(dst + f) = ld_sys_v4(src + (f - fb));
    } else {
      for (int j = 0; j < 4 && f + j < total; ++j) {
        const uint32_t key = v.begin[0] + static_cast<uint32_t>((f + j) / k);
        const int sj = owner_of(v, key);
        dst[f + j] = ld_sys(static_cast<const float*>(v.weights[sj]) + (f + j - flat_begin(v, sj)));
      }
    }
  }
}

__global__ void __launch_bounds__(256)
push_dense_f32_kernel(const __grid_constant__ MpsTableView v, int rot_shard, float* __restrict__ src,
                      int zero_src) {
  const int S = v.num_shards;
  const int64_t total = flat_begin(v, S);
  const int64_t rot = flat_begin(v, rot_shard) & ~int64_t(3);
  const int64_t nchunk = (total + 3) >> 2;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  const int k = v.vals_per_key;
  for (int64_t ci = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; ci < nchunk; ci += stride) {
    int64_t f = (ci << 2) + rot;
    if (f >= (nchunk << 2)) f -= (nchunk << 2);
    if (f >= total) continue;
    const uint32_t key0 = v.begin[0] + static_cast<uint32_t>(f / k);
    const int s = owner_of(v, key0);
    const int64_t fb = flat_begin(v, s), fe = flat_begin(v, s + 1);
    float* base = static_cast<float*>(v.staging[s] ? v.staging[s] : v.weights[s]);
    if (f + 4 <= fe && ((f - fb) & 3) == 0) {
      const float4 x = *reinterpret_cast<const float4*>(src + f);
      if (x.x != 0.f || x.y != 0.f || x.z != 0.f || x.w != 0.f) {
        red_sys_v4(base + (f - fb), x);
        if (zero_src) *reinterpret_cast<float4*>(src + f) = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    } else {
      for (int j = 0; j < 4 && f + j < total; ++j) {
        const float x = src[f + j];
        if (x == 0.f) continue;
        const uint32_t key = v.begin[0] + static_cast<uint32_t>((f + j) / k);
        const int sj = owner_of(v, key);
        float* bj = static_cast<float*>(v.staging[sj] ? v.staging[sj] : v.weights[sj]);
        red_sys(bj + (f + j - flat_begin(v, sj)), x);
        if (zero_src) src[f + j] = 0.f;
      }
    }
  }
}

// f64 tables (reference Val = double, worker/kv_client_table.hpp:98-99): scalar path.
__global__ void __launch_bounds__(256)
pull_dense_f64_kernel(const __grid_constant__ MpsTableView v, const __grid_constant__ MpsClockView c,
                      int gated, int progress, double* __restrict__ dst) {
  if (gated) cta_gate(c, progress);
  const int64_t total = flat_begin(v, v.num_shards);
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  const int k = v.vals_per_key;
  for (int64_t f = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; f < total; f += stride) {
    const int s = owner_of(v, v.begin[0] + static_cast<uint32_t>(f / k));
    dst[f] = ld_sys(static_cast<const double*>(v.weights[s]) + (f - flat_begin(v, s)));
  }
}

__global__ void __launch_bounds__(256)
push_dense_f64_kernel(const __grid_constant__ MpsTableView v, double* __restrict__ src, int zero_src) {
  const int64_t total = flat_begin(v, v.num_shards);
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  const int k = v.vals_per_key;
  for (int64_t f = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; f < total; f += stride) {
    const double x = src[f];
    if (x == 0.0) continue;
    const int s = owner_of(v, v.begin[0] + static_cast<uint32_t>(f / k));
    double* base = static_cast<double*>(v.staging[s] ? v.staging[s] : v.weights[s]);
    red_sys(base + (f - flat_begin(v, s)), x);
    if (zero_src) src[f] = 0.0;
  }
}

static int grid_for(int64_t n, int threads, int per_thread) {
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  int64_t want = (n + static_cast<int64_t>(threads) * per_thread - 1) / (static_cast<int64_t>(threads) * per_thread);
  int64_t cap = static_cast<int64_t>(sms) * 8;  // 8 x 256 threads = full occupancy per SM
  if (want < 1) want = 1;
  return static_cast<int>(want < cap ? want : cap);
}

static MpsClockView null_clock() {
  MpsClockView c{};
  return c;
}

}  // namespace mps

extern "C" {

int mps_pull(const MpsTableView* view, const MpsClockView* clk, int worker, int progress,
             const uint32_t* keys, void* vals, int64_t n, void* stream) {
  (void)worker;
  if (n <= 0) return 0;
  auto st = static_cast<cudaStream_t>(stream);
  const int grid = mps::grid_for(n, 256, 2);
  const MpsClockView c = clk ? *clk : mps::null_clock();
  if (view->dtype == MPS_F64)
    mps::pull_kernel<double><<<grid, 256, 0, st>>>(*view, c, clk != nullptr, progress, keys,
                                                   static_cast<double*>(vals), n);
  else
    mps::pull_kernel<float><<<grid, 256, 0, st>>>(*view, c, clk != nullptr, progress, keys,
                                                  static_cast<float*>(vals), n);
  MPS_LAUNCH_CHECK();
  return 0;
}

int mps_push(const MpsTableView* view, const uint32_t* keys, const void* vals, int64_t n, void* stream) {
  if (n <= 0) return 0;
  auto st = static_cast<cudaStream_t>(stream);
  const int grid = mps::grid_for(n, 256, 2);
  if (view->dtype == MPS_F64)
    mps::push_kernel<double><<<grid, 256, 0, st>>>(*view, keys, static_cast<const double*>(vals), n);
  else
    mps::push_kernel<float><<<grid, 256, 0, st>>>(*view, keys, static_cast<const float*>(vals), n);
  MPS_LAUNCH_CHECK();
  return 0;
}

int mps_pull_dense(const MpsTableView* view, const MpsClockView* clk, int worker, int progress,
                   void* dst, void* stream) {
  (void)worker;
  auto st = static_cast<cudaStream_t>(stream);
  const int64_t total = static_cast<int64_t>(view->begin[view->num_shards] - view->begin[0]) * view->vals_per_key;
  if (total <= 0) return 0;
  const MpsClockView c = clk ? *clk : mps::null_clock();
  const int rot = clk ? clk->own_first : 0;
  if (view->dtype == MPS_F64) {
    mps::pull_dense_f64_kernel<<<mps::grid_for(total, 256, 2), 256, 0, st>>>(*view, c, clk != nullptr, progress,
                                                                             static_cast<double*>(dst));
  } else {
    mps::pull_dense_f32_kernel<<<mps::grid_for(total / 4 + 1, 256, 2), 256, 0, st>>>(
        *view, c, clk != nullptr, progress, rot, static_cast<float*>(dst));
  }
  MPS_LAUNCH_CHECK();
  return 0;
}

int mps_push_dense(const MpsTableView* view, void* src, int zero_src, int rot_shard, void* stream) {
  auto st = static_cast<cudaStream_t>(stream);
  const int64_t total = static_cast<int64_t>(view->begin[view->num_shards] - view->begin[0]) * view->vals_per_key;
  if (total <= 0) return 0;
  if (view->dtype == MPS_F64) {
    mps::push_dense_f64_kernel<<<mps::grid_for(total, 256, 2), 256, 0, st>>>(
        *view, static_cast<double*>(src), zero_src);
  } else {
    mps::push_dense_f32_kernel<<<mps::grid_for(total / 4 + 1, 256, 2), 256, 0, st>>>(
        *view, rot_shard, static_cast<float*>(src), zero_src);
  }
  MPS_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
