// probe.cu — NVLink store / fence micro-measurements used to choose the exchange protocol of the LR
// kernel (tools/nvlink_probe.py prints them; numbers in profiles/nvlink_probe_*.txt).
//
// One launch on ONE rank (nothing waits on a peer): every thread stores its share of `bytes` into a
// destination (a peer's buffer, this GPU's buffer, or the multicast mapping), the CTA meets on
// bar.sync, thread 0 takes the time, executes fence.acq_rel.sys and takes the time again.
//   out[0] = max over CTAs of (stores issued - start)   ns
//   out[1] = max over CTAs of (fence done  - start)     ns   = when a flag could be published
//   out[2] = max over CTAs of the fence alone           ns
#include "lr_common.cuh"

namespace mps {

enum { kProbeSt16 = 0, kProbeMcSt16 = 1, kProbeBulk = 2, kProbeRed16 = 3 };

__global__ void __launch_bounds__(512, 2)
probe_store_kernel(int mode, float* __restrict__ dst, const float* __restrict__ src, int64_t n_floats,
                   unsigned long long* __restrict__ out) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int64_t tid = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int64_t nthreads = static_cast<int64_t>(gridDim.x) * blockDim.x;
  __syncthreads();
  const unsigned long long t0 = globaltimer_ns();
  if (mode == kProbeBulk) {
    // every CTA owns a contiguous slice; it stages 16 KiB pieces in shared memory and hands them to the
    // copy engine (cp.async.bulk shared -> global), which writes full lines over NVLink
    constexpr int kPiece = 16384 / 4;
    const int64_t per = (n_floats / gridDim.x) & ~int64_t(3);
    const int64_t b0 = per * blockIdx.x;
    float* sm = reinterpret_cast<float*>(smem);
    for (int64_t off = 0; off < per; off += kPiece) {
      const int64_t len = per - off < kPiece ? per - off : kPiece;
      if (off > 0) {
        if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        __syncthreads();
      }
      for (int64_t i = threadIdx.x * 4; i < len; i += blockDim.x * 4)
        *reinterpret_cast<float4*>(sm + i) = *reinterpret_cast<const float4*>(src + b0 + off + i);
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncthreads();
      if (threadIdx.x == 0) {
        const uint32_t s = static_cast<uint32_t>(__cvta_generic_to_shared(sm));
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst + b0 + off), "r"(s),
                     "r"(static_cast<uint32_t>(len * 4))
                     : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
    }
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  } else {
    const int64_t nc = n_floats / 4;
    for (int64_t i = tid; i < nc; i += nthreads) {
      Chunk<float> c;
      c.v = *reinterpret_cast<const float4*>(src + 4 * i);
      if (mode == kProbeSt16) st_sys_chunk(dst + 4 * i, c);
      else if (mode == kProbeMcSt16) mc_st_chunk(dst + 4 * i, c);
      else red_sys_v4(dst + 4 * i, c.v);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned long long t1 = globaltimer_ns();
    fence_sys();
    const unsigned long long t2 = globaltimer_ns();
    atomicMax(out + 0, t1 - t0);
    atomicMax(out + 1, t2 - t0);
    atomicMax(out + 2, t2 - t1);
  }
}

}  // namespace mps

extern "C" int mps_probe_store(int mode, void* dst, const void* src, int64_t bytes, unsigned long long* out,
                               void* stream) {
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  auto st = static_cast<cudaStream_t>(stream);
  MPS_CHECK(cudaMemsetAsync(out, 0, 3 * sizeof(unsigned long long), st));
  MPS_CHECK(cudaFuncSetAttribute(mps::probe_store_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384));
  mps::probe_store_kernel<<<2 * sms, 512, 16384, st>>>(mode, static_cast<float*>(dst), static_cast<const float*>(src),
                                                       bytes / 4, out);
  MPS_LAUNCH_CHECK();
  return 0;
}
