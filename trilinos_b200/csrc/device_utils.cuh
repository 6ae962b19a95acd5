// Small device helpers shared by the kernels (sm_100a).
#ifndef B200_DEVICE_UTILS_CUH
#define B200_DEVICE_UTILS_CUH

#include <cuda_runtime.h>

#include <cstdint>

namespace b200 {

// Streaming 256-bit load (sm_100 LDG.256): read-only path, no L1 allocation, evict-first in
// L2 — each Krylov-basis element is touched exactly once per pass, so it must not push the
// small re-used operand (the n x b block, which fits the 126 MB L2) out of cache.
struct double4v { double x, y, z, w; };
__device__ __forceinline__ double4v ld_stream4(const double* p) {
  double4v v;
  asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ double ld_stream1(const double* p) {
  double v;
  asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}
// re-used operand: 256-bit load that asks L2 to keep the line
__device__ __forceinline__ double4v ld_keep4(const double* p) {
  double4v v;
  asm volatile("ld.global.nc.L2::evict_last.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w)
               : "l"(p));
  return v;
}
// plain (coherent) 256-bit load / store for operands that are also written by the kernel
__device__ __forceinline__ double4v ld_plain4(const double* p) {
  double4v v;
  asm volatile("ld.global.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ void st_plain4(double* p, const double4v& v) {
  asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(v.x), "d"(v.y), "d"(v.z),
               "d"(v.w)
               : "memory");
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Publish protocol for host-visible results: a thread that stored results into the mapped
// mailbox calls fence_results() (system-scope fence, ~3 us if a whole CTA does it, so only
// the writers do), the CTA synchronises, then one thread releases the sequence number the
// host is spinning on.
__device__ __forceinline__ void fence_results() { __threadfence_system(); }
__device__ __forceinline__ void publish_flag(uint64_t* flag, uint64_t seq) {
  __syncthreads();
  if (flag != nullptr && threadIdx.x == 0) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(flag), "l"(seq) : "memory");
  }
}

// Last-CTA election for two-stage deterministic reductions.  Returns true in every thread of
// the CTA that arrived last; the counter resets itself for the next launch.
__device__ __forceinline__ bool last_cta_arrives(unsigned int* ticket, unsigned int total) {
  __shared__ bool is_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int t = atomicAdd(ticket, 1u);
    is_last = (t == total - 1);
    if (is_last) *ticket = 0u;
  }
  __syncthreads();
  if (is_last) __threadfence();
  return is_last;
}

}  // namespace b200
#endif
