// FP64 issue-rate probe (B200): DFMA (SIMT) vs DMMA (mma.sync.m8n8k4.f64) FMA/clk/SM.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/dmma_probe.cu -o tools/dmma_probe && tools/dmma_probe
#include <cstdio>
#include <cuda_runtime.h>

__global__ void dfma_kernel(double* out, int iters) {
  double a = threadIdx.x * 1e-9, b = 1.0000001;
  double acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = fma(acc[i], b, a);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void dmma_kernel(double* out, int iters) {
  double a = threadIdx.x * 1e-9, b = 1.0000001;
  double c[8][2];
#pragma unroll
  for (int i = 0; i < 8; ++i) c[i][0] = c[i][1] = i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                   : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
  int sms = 0, khz = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  double* out;
  cudaMalloc(&out, sizeof(double) * sms * 4 * 512);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 20000;
  for (int which = 0; which < 2; ++which)
    for (int warps = 4; warps <= 16; warps *= 2) {
      for (int rep = 0; rep < 2; ++rep) {
        cudaEventRecord(e0);
        if (which == 0) dfma_kernel<<<sms, warps * 32>>>(out, iters);
        else dmma_kernel<<<sms, warps * 32>>>(out, iters);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
      }
      float ms = 0;
      cudaEventElapsedTime(&ms, e0, e1);
      const double fmas = which == 0 ? (double)sms * warps * 32 * 16 * iters : (double)sms * warps * 8 * 256.0 * iters;
      printf("%s warps/SM=%2d: %.2f ms  %.2f TFLOP/s  %.1f FMA/clk/SM (at %d MHz nominal)\n", which ? "DMMA m8n8k4" : "DFMA       ",
             warps, ms, 2 * fmas / ms * 1e-9, fmas / (ms * 1e-3) / sms / (khz * 1e3), khz / 1000);
    }
  return 0;
}
