// hash_table.cu — MapStorage on device: a peer-reachable open-addressing hash shard.
//
// Reference server/map_storage.hpp:13-26: std::map<Key,Val>, `+=` on add, operator[] (inserting
// a zero) on get.  Device form: every shard is [capacity] u32 keys (0xFFFFFFFF = empty) and
// [capacity] f32 values inside the owner's symmetric slab; any rank inserts with
// atom.cas.sys on the key slot and accumulates with red.add.sys on the value, so the table
// grows only with the keys actually touched (the point of MapStorage: sparse, unbounded use
// of a huge key range).  Linear probing; capacity is a power of two.
#include "common.cuh"

namespace mps {

constexpr uint32_t kEmptyKey = 0xFFFFFFFFu;

__device__ __forceinline__ uint32_t hash32(uint32_t x) {
  x ^= x >> 16;
  x *= 0x7feb352du;
  x ^= x >> 15;
  x *= 0x846ca68bu;
  x ^= x >> 16;
  return x;
}
__device__ __forceinline__ uint32_t cas_sys(uint32_t* p, uint32_t expected, uint32_t desired) {
  uint32_t old;
  asm volatile("atom.relaxed.sys.global.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(p), "r"(expected), "r"(desired)
      : "memory");
  return old;
}
__device__ __forceinline__ uint32_t ld_key_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// returns the slot of `key` in the shard, inserting it if absent; -1 when the shard is full
__device__ __forceinline__ int64_t find_or_insert(uint32_t* keys, uint32_t cap_mask, uint32_t key) {
  uint32_t h = hash32(key) & cap_mask;
  for (uint32_t probe = 0; probe <= cap_mask; ++probe) {
    uint32_t cur = ld_key_sys(keys + h);
    if (cur == key) return h;
    if (cur == kEmptyKey) {
      cur = cas_sys(keys + h, kEmptyKey, key);
      if (cur == kEmptyKey || cur == key) return h;
    }
    h = (h + 1) & cap_mask;
  }
  return -1;
}
__device__ __forceinline__ int64_t find_only(const uint32_t* keys, uint32_t cap_mask, uint32_t key) {
  uint32_t h = hash32(key) & cap_mask;
  for (uint32_t probe = 0; probe <= cap_mask; ++probe) {
    const uint32_t cur = ld_key_sys(keys + h);
    if (cur == key) return h;
    if (cur == kEmptyKey) return -1;
    h = (h + 1) & cap_mask;
  }
  return -1;
}

__global__ void __launch_bounds__(256)
hash_add_kernel(const __grid_constant__ MpsHashView v, const uint32_t* __restrict__ keys,
    const float* __restrict__ vals,
                int64_t n, int* __restrict__ overflow) {
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint32_t key = keys[i];
    int s = -1;
    for (int j = 0; j < v.num_shards; ++j)
      if (key >= v.begin[j] && key < v.begin[j + 1]) { s = j; break; }
    if (s < 0) continue;  // out-of-range keys are dropped, like RangePartitionManager::Slice
    const int64_t slot = find_or_insert(static_cast<uint32_t*>(v.keys[s]), v.capacity[s] - 1, key);
    if (slot < 0) { atomicExch(overflow, 1); continue; }
    red_sys(static_cast<float*>(v.vals[s]) + slot, vals[i]);
  }
}

__global__ void __launch_bounds__(256)
hash_get_kernel(const __grid_constant__ MpsHashView v, const __grid_constant__ MpsClockView c, int gated, int progress,
                const uint32_t* __restrict__ keys, float* __restrict__ vals, int64_t n) {
  if (gated) cta_gate(c, progress);
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint32_t key = keys[i];
    int s = -1;
    for (int j = 0; j < v.num_shards; ++j)
      if (key >= v.begin[j] && key < v.begin[j + 1]) { s = j; break; }
    float out = 0.f;
    if (s >= 0) {
      const int64_t slot = find_only(static_cast<const uint32_t*>(v.keys[s]), v.capacity[s] - 1, key);
      if (slot >= 0) out = ld_sys(static_cast<const float*>(v.vals[s]) + slot);
    }
    vals[i] = out;
  }
}

__global__ void hash_clear_kernel(uint32_t* keys, float* vals, int64_t cap) {
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < cap; i += stride) {
    keys[i] = kEmptyKey;
    vals[i] = 0.f;
  }
}

// compacts the live (key, value) pairs of one LOCAL shard (checkpoint dump / size)
__global__ void hash_export_kernel(const uint32_t* __restrict__ keys, const float* __restrict__ vals, int64_t cap,
                                   uint32_t* __restrict__ out_keys, float* __restrict__ out_vals,
                                   int* __restrict__ count) {
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < cap; i += stride) {
    const uint32_t k = keys[i];
    if (k != kEmptyKey) {
      const int pos = atomicAdd(count, 1);
      out_keys[pos] = k;
      out_vals[pos] = vals[i];
    }
  }
}

}  // namespace mps

extern "C" {

int mps_hash_clear(void* keys, void* vals, int64_t capacity, void* stream) {
  mps::hash_clear_kernel<<<296, 256, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<uint32_t*>(keys),
                                                                          static_cast<float*>(vals), capacity);
  MPS_LAUNCH_CHECK();
  return 0;
}

int mps_hash_add(const MpsHashView* view, const uint32_t* keys, const float* vals, int64_t n, int* overflow_flag,
                 void* stream) {
  if (n <= 0) return 0;
  const int grid = static_cast<int>((n + 255) / 256 < 1184 ? (n + 255) / 256 : 1184);
  mps::hash_add_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(*view, keys, vals, n, overflow_flag);
  MPS_LAUNCH_CHECK();
  return 0;
}

int mps_hash_get(const MpsHashView* view, const MpsClockView* clk, int progress, const uint32_t* keys, float* vals,
                 int64_t n, void* stream) {
  if (n <= 0) return 0;
  const int grid = static_cast<int>((n + 255) / 256 < 1184 ? (n + 255) / 256 : 1184);
  MpsClockView c{};
  if (clk) c = *clk;
  mps::hash_get_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(*view, c, clk != nullptr, progress, keys,
                                                                            vals, n);
  MPS_LAUNCH_CHECK();
  return 0;
}

int mps_hash_export(const void* keys, const void* vals, int64_t capacity, uint32_t* out_keys, float* out_vals,
                    int* count, void* stream) {
  auto st = static_cast<cudaStream_t>(stream);
  MPS_CHECK(cudaMemsetAsync(count, 0, sizeof(int), st));
  mps::hash_export_kernel<<<296, 256, 0, st>>>(static_cast<const uint32_t*>(keys), static_cast<const float*>(vals),
                                               capacity, out_keys, out_vals, count);
  MPS_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
