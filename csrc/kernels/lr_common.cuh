// lr_common.cuh — pieces shared by the sparse-LR kernels (lr.cu, lr_xchg.cu), templated on the
// table's value type T: float, or double for the reference's Val (every reference app runs
// KVClientTable<double>, apps/logistic_regression/lr_example.cpp:182,242).  Feature values and
// labels of the samples stay fp32 in memory (they are inputs, widened exactly on load); weights,
// logits, sigmoid, gradients and every accumulator are T.
#pragma once
#include "common.cuh"

namespace mps {

// ---------------------------------------------------------------- 16-byte chunks of T
template <typename T> struct Chunk;
template <> struct Chunk<float> {
  static constexpr int N = 4;
  float4 v;
  __device__ __forceinline__ float get(int i) const { return i == 0 ? v.x : (i == 1 ? v.y : (i == 2 ? v.z : v.w)); }
  __device__ __forceinline__ bool any() const { return v.x != 0.f || v.y != 0.f || v.z != 0.f || v.w != 0.f; }
  __device__ __forceinline__ void add(const Chunk& o) { v.x += o.v.x; v.y += o.v.y; v.z += o.v.z; v.w += o.v.w; }
  static __device__ __forceinline__ Chunk zero() { Chunk c; c.v = make_float4(0.f, 0.f, 0.f, 0.f); return c; }
};
template <> struct Chunk<double> {
  static constexpr int N = 2;
  double2 v;
  __device__ __forceinline__ double get(int i) const { return i == 0 ? v.x : v.y; }
  __device__ __forceinline__ bool any() const { return v.x != 0.0 || v.y != 0.0; }
  __device__ __forceinline__ void add(const Chunk& o) { v.x += o.v.x; v.y += o.v.y; }
  static __device__ __forceinline__ Chunk zero() { Chunk c; c.v = make_double2(0.0, 0.0); return c; }
};

// strong system-scope chunk load (peer shard or peer scratch)
__device__ __forceinline__ Chunk<float> ld_sys_chunk(const float* p) { Chunk<float> c; c.v = ld_sys_v4(p); return c; }
__device__ __forceinline__ Chunk<double> ld_sys_chunk(const double* p) {
  Chunk<double> c;
  c.v = ld_sys_v2(p);
  return c;
}
// L2 (cache-global) chunk load of local memory other CTAs have been adding into
__device__ __forceinline__ Chunk<float> ld_cg_chunk(const float* p) { Chunk<float> c; c.v = ld_cg_v4(p); return c; }
__device__ __forceinline__ Chunk<double> ld_cg_chunk(const double* p) {
  Chunk<double> c;
  asm volatile("ld.global.cg.v2.f64 {%0,%1}, [%2];" : "=d"(c.v.x), "=d"(c.v.y) : "l"(p) : "memory");
  return c;
}
__device__ __forceinline__ double ld_cg(const double* p) {
  double v;
  asm volatile("ld.global.cg.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_chunk(float* p, const Chunk<float>& c) { *reinterpret_cast<float4*>(p) = c.v; }
__device__ __forceinline__ void st_chunk(double* p, const Chunk<double>& c) { *reinterpret_cast<double2*>(p) = c.v; }
// system-scope store into a peer's memory
__device__ __forceinline__ void st_sys_chunk(float* p, const Chunk<float>& c) {
  asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(c.v.x), "f"(c.v.y), "f"(c.v.z),
               "f"(c.v.w)
               : "memory");
}
__device__ __forceinline__ void st_sys_chunk(double* p, const Chunk<double>& c) {
  asm volatile("st.relaxed.sys.global.v2.f64 [%0], {%1,%2};" ::"l"(p), "d"(c.v.x), "d"(c.v.y) : "memory");
}
__device__ __forceinline__ void st_sys(float* p, float x) {
  asm volatile("st.relaxed.sys.global.f32 [%0], %1;" ::"l"(p), "f"(x) : "memory");
}
__device__ __forceinline__ void st_sys(double* p, double x) {
  asm volatile("st.relaxed.sys.global.f64 [%0], %1;" ::"l"(p), "d"(x) : "memory");
}
// vector red into a (peer) shard: one 16-byte packet for fp32; fp64 has no vector red
__device__ __forceinline__ void red_sys_chunk(float* p, const Chunk<float>& c) { red_sys_v4(p, c.v); }
__device__ __forceinline__ void red_sys_chunk(double* p, const Chunk<double>& c) {
  if (c.v.x != 0.0) red_sys(p, c.v.x);
  if (c.v.y != 0.0) red_sys(p + 1, c.v.y);
}
__device__ __forceinline__ void red_gpu(double* p, double v) {
  asm volatile("red.relaxed.gpu.global.add.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
}

// NVLS: one load that returns the sum over every GPU's copy (reduced in the switch), one store that
// lands in every GPU's copy.  `p` is an address inside the multicast mapping of a symmetric buffer.
__device__ __forceinline__ Chunk<float> mc_ld_reduce_chunk(const float* p) {
  Chunk<float> c;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(c.v.x), "=f"(c.v.y), "=f"(c.v.z), "=f"(c.v.w)
               : "l"(p)
               : "memory");
  return c;
}
__device__ __forceinline__ Chunk<double> mc_ld_reduce_chunk(const double* p) {
  Chunk<double> c;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f64 %0, [%1];" : "=d"(c.v.x) : "l"(p) : "memory");
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f64 %0, [%1];" : "=d"(c.v.y) : "l"(p + 1) : "memory");
  return c;
}
__device__ __forceinline__ float mc_ld_reduce(const float* p) {
  float v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ double mc_ld_reduce(const double* p) {
  double v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void mc_st_chunk(float* p, const Chunk<float>& c) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(c.v.x), "f"(c.v.y),
               "f"(c.v.z), "f"(c.v.w)
               : "memory");
}
// a store does not interpret its payload: the two doubles travel as ONE 16-byte packet (multimem.st has no
// v2.f64 form; two scalar f64 stores make twice the packets and drain slowly)
__device__ __forceinline__ void mc_st_chunk(double* p, const Chunk<double>& c) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p),
               "f"(__int_as_float(__double2loint(c.v.x))), "f"(__int_as_float(__double2hiint(c.v.x))),
               "f"(__int_as_float(__double2loint(c.v.y))), "f"(__int_as_float(__double2hiint(c.v.y)))
               : "memory");
}
__device__ __forceinline__ void mc_st(float* p, float x) {
  asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(p), "f"(x) : "memory");
}
__device__ __forceinline__ void mc_st(double* p, double x) {
  asm volatile("multimem.st.relaxed.sys.global.f64 [%0], %1;" ::"l"(p), "d"(x) : "memory");
}

// ---------------------------------------------------------------- LR math in T
__device__ __forceinline__ float sigmoid_t(float z) { return 1.f / (1.f + __expf(-z)); }
__device__ __forceinline__ double sigmoid_t(double z) { return 1.0 / (1.0 + exp(-z)); }
// logloss of label y in {0,1} at logit z, numerically stable: softplus(z) - y*z
__device__ __forceinline__ float logloss_t(float z, float y) {
  return fmaxf(z, 0.f) + log1pf(__expf(-fabsf(z))) - y * z;
}
__device__ __forceinline__ double logloss_t(double z, double y) {
  return fmax(z, 0.0) + log1p(exp(-fabs(z))) - y * z;
}

// block reduce of (loss, correct) pairs then one atomic pair per CTA
__device__ __forceinline__ void block_accumulate(float loss, float correct, float* out) {
  __shared__ float sl[32], sc[32];
  loss = warp_sum(loss);
  correct = warp_sum(correct);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) { sl[wid] = loss; sc[wid] = correct; }
  __syncthreads();
  if (wid == 0) {
    const int nw = (blockDim.x + 31) >> 5;
    float a = lane < nw ? sl[lane] : 0.f, b = lane < nw ? sc[lane] : 0.f;
    a = warp_sum(a);
    b = warp_sum(b);
    if (lane == 0 && (a != 0.f || b != 0.f)) {
      atomicAdd(out, a);
      atomicAdd(out + 1, b);
    }
  }
  __syncthreads();
}

template <int L, typename T>
__device__ __forceinline__ T group_sum(T v, unsigned mask) {
#pragma unroll
  for (int o = L / 2; o > 0; o >>= 1) v += __shfl_xor_sync(mask, v, o);
  return v;
}

// flat index of shard s's first key (one value per key in the LR tables)
__device__ __forceinline__ int64_t flat_begin1(const MpsTableView& v, int s) {
  return static_cast<int64_t>(v.begin[s] - v.begin[0]);
}

// every interior shard boundary on a multiple of the chunk width?
template <typename T>
__device__ __forceinline__ bool shards_chunk_aligned(const MpsTableView& v) {
  bool ok = true;
  for (int s = 1; s < v.num_shards; ++s) ok = ok && ((flat_begin1(v, s) & (Chunk<T>::N - 1)) == 0);
  return ok;
}

// address, valid on this device, of flat element f inside its owning shard
template <typename T>
__device__ __forceinline__ T* shard_elem(const MpsTableView& v, int64_t f) {
  const int s = owner_of(v, v.begin[0] + static_cast<uint32_t>(f));
  return static_cast<T*>(v.weights[s]) + (f - flat_begin1(v, s));
}

// Rows of a batch against a local dense copy of the weights.  L lanes cooperate on one row
// (32/L rows in flight per warp), each lane keeps up to E of the row's non-zeros in
// registers so the weight gathers of a row are all issued before the first one is used.
// kTrain: also scatter alpha*x*(y-pred) into gacc (red into the local L2; kSysRed: at system
// scope, for accumulators that peers read afterwards).
// Feature ids are key ids: wloc / gacc are indexed by (id - key0); ids outside
// [key0, key0 + n_keys) are skipped, like the reference's Slice drops out-of-range keys
// (base/range_partition_manager.hpp:24-41).
constexpr int kRowRegs = 8;

template <typename T>
__device__ __forceinline__ void red_local(T* p, T v, bool sys) {
  if (sys) red_sys(p, v); else red_gpu(p, v);
}

// The two sample layouts a kernel can read (MpsCsr.format): plain CSR arrays, or the compact form
// the host->device loader ships (u8 row lengths + an offset every 8 rows, bit-packed feature ids, u8
// labels) decoded on the fly, so an end-to-end step needs no separate unpack pass.
template <bool kPacked> struct CsrView;
template <> struct CsrView<false> {
  const MpsCsr& m;
  template <int L>
  __device__ __forceinline__ void row(int64_t ridx, int, unsigned, int64_t& beg, int64_t& end) const {
    beg = m.row_ptr[ridx];
    end = m.row_ptr[ridx + 1];
  }
  __device__ __forceinline__ uint32_t col(int64_t j) const { return ld_stream(m.col + j); }
  __device__ __forceinline__ float label(int64_t r) const { return m.label[r]; }
};
template <> struct CsrView<true> {
  const MpsCsr& m;
  // offset of row ridx = offset of its block of 8 rows + the lengths of the rows before it in the
  // block, summed by the row's own lanes (L >= 8 lanes per row, all of them call this together)
  template <int L>
  __device__ __forceinline__ void row(int64_t ridx, int gl, unsigned mask, int64_t& beg, int64_t& end) const {
    const int k = static_cast<int>(ridx & 7);
    int part = (gl < k) ? static_cast<int>(m.row_len[(ridx & ~int64_t(7)) + gl]) : 0;
#pragma unroll
    for (int o = L / 2; o > 0; o >>= 1) part += __shfl_xor_sync(mask, part, o);
    beg = m.coarse[ridx >> 3] + part;
    end = beg + m.row_len[ridx];
  }
  __device__ __forceinline__ uint32_t col(int64_t j) const {
    const int64_t bit = m.bit_off + j * m.id_bits;
    const uint32_t* w = static_cast<const uint32_t*>(m.packed_ids) + (bit >> 5);
    const uint64_t two = (static_cast<uint64_t>(ld_stream(w + 1)) << 32) | ld_stream(w);
    return static_cast<uint32_t>(two >> (bit & 31)) & ((1u << m.id_bits) - 1u);
  }
  __device__ __forceinline__ float label(int64_t r) const { return static_cast<float>(m.label_u8[r]); }
};

template <int L, bool kTrain, typename T, bool kPacked = false>
__device__ __forceinline__ void rows_pass(const MpsCsr& csr, int64_t row0, int64_t batch, uint32_t key0,
                                          uint32_t n_keys, const T* __restrict__ wloc, T* __restrict__ gacc,
                                          T alpha, bool sys_red, int64_t group, int64_t ngroups, float& loss,
                                          float& correct) {
  constexpr int E = kRowRegs;
  const CsrView<kPacked> view{csr};
  const int lane = threadIdx.x & 31;
  const int gl = lane & (L - 1);
  const unsigned mask = (L == 32) ? 0xffffffffu : (((1u << L) - 1u) << (lane & ~(L - 1)));
  for (int64_t r = group; r < batch; r += ngroups) {
    int64_t ridx = row0 + r;
    if (ridx >= csr.n_rows) ridx %= csr.n_rows;
    int64_t beg, end;
    view.template row<L>(ridx, gl, mask, beg, end);
    uint32_t c[E];
    T x[E];
#pragma unroll
    for (int e = 0; e < E; ++e) {
      const int64_t j = beg + gl + e * L;
      c[e] = 0xffffffffu;
      x[e] = T(0);
      if (j < end) {
        c[e] = view.col(j) - key0;
        x[e] = static_cast<T>(ld_stream(csr.val + j));
        if (c[e] >= n_keys) x[e] = T(0);
      }
    }
    T w[E];
#pragma unroll
    for (int e = 0; e < E; ++e) w[e] = (c[e] < n_keys) ? ld_cg(wloc + c[e]) : T(0);
    T acc = T(0);
#pragma unroll
    for (int e = 0; e < E; ++e) acc = fma(w[e], x[e], acc);
    for (int64_t j = beg + gl + E * L; j < end; j += L) {  // rows longer than L*E non-zeros
      const uint32_t cj = view.col(j) - key0;
      if (cj < n_keys) acc = fma(ld_cg(wloc + cj), static_cast<T>(ld_stream(csr.val + j)), acc);
    }
    const T z = group_sum<L>(acc, mask);
    const T y = static_cast<T>(view.label(ridx));
    const T pred = sigmoid_t(z);
    if (gl == 0) {
      loss += static_cast<float>(logloss_t(z, y));
      correct += ((pred > T(0.5)) == (y > T(0.5))) ? 1.f : 0.f;
    }
    if (kTrain) {
      const T g = alpha * (y - pred);
#pragma unroll
      for (int e = 0; e < E; ++e)
        if (x[e] != T(0)) red_local(gacc + c[e], g * x[e], sys_red);
      for (int64_t j = beg + gl + E * L; j < end; j += L) {
        const T kx = static_cast<T>(ld_stream(csr.val + j));
        const uint32_t cj = view.col(j) - key0;
        if (kx != T(0) && cj < n_keys) red_local(gacc + cj, g * kx, sys_red);
      }
    }
  }
}

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// lanes per row from the average row length: 8 lanes hold rows of <= 64 nnz in registers
static inline int lanes_for(int64_t avg_nnz) {
  if (avg_nnz <= 8 * kRowRegs) return 8;
  if (avg_nnz <= 16 * kRowRegs) return 16;
  return 32;
}

}  // namespace mps
