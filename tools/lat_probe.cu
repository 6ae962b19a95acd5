// latency probe: host-visible round trip of tiny kernels on B200 (build: nvcc -arch=sm_100a)
#include <cuda_runtime.h>
#include <chrono>
#include <cstdio>
#include <cstdint>
#include <immintrin.h>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("err %s line %d\n", cudaGetErrorString(e), __LINE__); return 1;}}while(0)
__global__ void k_empty() {}
__global__ void k_flag(uint64_t* flag, uint64_t seq) {
  if (threadIdx.x == 0) asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(flag), "l"(seq) : "memory");
}
__global__ void k_flag_fence(double* out, uint64_t* flag, uint64_t seq) {
  if (threadIdx.x < 4) out[threadIdx.x] = 1.0 * seq;
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(flag), "l"(seq) : "memory");
}
// many CTAs: partial sums + ticket + last CTA publishes
__global__ void k_reduce(const double* x, int64_t n, double* ws, unsigned* ticket, double* out, uint64_t* flag, uint64_t seq, int mode) {
  double acc = 0;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) acc += x[i] * x[i];
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  __shared__ double sm[8]; __shared__ bool last;
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) { double t = 0; for (int w = 0; w < 8; ++w) t += sm[w]; ws[blockIdx.x] = t; __threadfence(); unsigned tk = atomicAdd(ticket, 1u); last = (tk == gridDim.x - 1); if (last) *ticket = 0; }
  __syncthreads();
  if (!last) return;
  if (threadIdx.x < 32) {
    double t = 0; volatile double* p = ws;
    for (int i = threadIdx.x; i < gridDim.x; i += 32) t += p[i];
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if (threadIdx.x == 0) { out[0] = t; if (mode == 0) { __threadfence_system(); asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(flag), "l"(seq) : "memory"); } }
  }
}
template <class F> double bench(const char* name, int reps, F f) {
  for (int i = 0; i < 20; ++i) f();
  auto t0 = std::chrono::steady_clock::now();
  for (int i = 0; i < reps; ++i) f();
  auto t1 = std::chrono::steady_clock::now();
  double us = std::chrono::duration<double, std::micro>(t1 - t0).count() / reps;
  printf("%-60s %8.2f us\n", name, us);
  return us;
}
int main() {
  cudaStream_t s; CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  void* hp; CK(cudaHostAlloc(&hp, 4096, cudaHostAllocMapped));
  volatile uint64_t* flag_h = (volatile uint64_t*)hp; *flag_h = 0;
  void* dp; CK(cudaHostGetDevicePointer(&dp, hp, 0));
  uint64_t* flag_d = (uint64_t*)dp; double* out_d = (double*)((char*)dp + 64);
  const int64_t n = 1000000; double* x; CK(cudaMalloc(&x, n * 8 * 4)); CK(cudaMemset(x, 0, n * 8 * 4));
  double* ws; CK(cudaMalloc(&ws, 1 << 20)); unsigned* ticket; CK(cudaMalloc(&ticket, 256)); CK(cudaMemset(ticket, 0, 256));
  double* dres; CK(cudaMalloc(&dres, 4096)); double* hres; CK(cudaHostAlloc((void**)&hres, 4096, 0));
  uint64_t seq = 0;
  auto wait = [&]() { while (*flag_h < seq) _mm_pause(); };
  bench("empty<<<1,32>>> + cudaStreamSynchronize", 2000, [&] { k_empty<<<1, 32, 0, s>>>(); cudaStreamSynchronize(s); });
  bench("flag<<<1,32>>> + host spin", 2000, [&] { ++seq; k_flag<<<1, 32, 0, s>>>(flag_d, seq); wait(); });
  bench("flag_fence<<<1,256>>> + host spin", 2000, [&] { ++seq; k_flag_fence<<<1, 256, 0, s>>>(out_d, flag_d, seq); wait(); });
  for (int grid : {64, 148, 296, 488, 1184}) {
    char nm[128]; snprintf(nm, 128, "reduce n=1e6 grid=%d, mapped publish + spin", grid);
    bench(nm, 2000, [&] { ++seq; k_reduce<<<grid, 256, 0, s>>>(x, n, ws, ticket, out_d, flag_d, seq, 0); wait(); });
  }
  bench("reduce grid=296 -> device, memcpyAsync D2H + streamSync", 2000, [&] { k_reduce<<<296, 256, 0, s>>>(x, n, ws, ticket, dres, flag_d, seq, 1); cudaMemcpyAsync(hres, dres, 8, cudaMemcpyDeviceToHost, s); cudaStreamSynchronize(s); });
  bench("reduce n=4e6 grid=592 mapped publish + spin", 2000, [&] { ++seq; k_reduce<<<592, 256, 0, s>>>(x, 4 * n, ws, ticket, out_d, flag_d, seq, 0); wait(); });
  // two dependent kernels then publish
  bench("empty + flag (2 launches) + spin", 2000, [&] { ++seq; k_empty<<<148, 256, 0, s>>>(); k_flag<<<1, 32, 0, s>>>(flag_d, seq); wait(); });
  bench("4x empty + flag (5 launches) + spin", 2000, [&] { ++seq; for (int i = 0; i < 4; ++i) k_empty<<<148, 256, 0, s>>>(); k_flag<<<1, 32, 0, s>>>(flag_d, seq); wait(); });
  // with event records in between (timing mode)
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  bench("event + reduce grid=296 + event + spin", 2000, [&] { ++seq; cudaEventRecord(e0, s); k_reduce<<<296, 256, 0, s>>>(x, n, ws, ticket, out_d, flag_d, seq, 0); cudaEventRecord(e1, s); wait(); });
  cudaStreamSynchronize(s); float ms; cudaEventElapsedTime(&ms, e0, e1); printf("last event-measured reduce: %.2f us\n", ms * 1e3);
  return 0;
}
