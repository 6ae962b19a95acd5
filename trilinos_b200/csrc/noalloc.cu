// b200_dot_noalloc: the column-wise dot products of MvDot for callers that must not allocate.
//
// No counterpart in the reference: added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The arithmetic is
// that of b200_mv_dot (reduce.cu; BelosEpetraAdapter.cpp:497-513 -> Epetra_MultiVector::Dot) on the rows this
// process holds: per-CTA partial sums, then ONE deterministic fold by the CTA that finishes last.  The other
// entry points keep those partials in the context's workspace (cudaMalloc at b200_ctx_create); here they live
// in a file-scope __device__ array of fixed size, so the entry point needs no context and allocates nothing:
// no cudaMalloc / cudaFree / cudaMallocAsync / pinned host memory, no library call that allocates.
//
// Limits (rejected with B200_ERR_ARG beyond them): ncols <= B200_NOALLOC_MAX_COLS (32); the grid is capped at
// kMaxCtas (1024) CTAs, rows are unbounded (grid-stride).  The array is kMaxCtas * 32 doubles + one 256-byte
// line for the ticket = 262400 bytes, held by every GPU that loads the library until the process exits.
//
// Every call on a device shares the array, which keeps the last call's contents: the partials are fully
// rewritten before they are read, the ticket is cleared on the caller's stream in front of every launch.
// Calls that overlap on the GPU (two streams, two host threads) race on it -- serialise them.
#include <cuda_runtime.h>

#include <mutex>

#include "b200_belos.h"
#include "device_utils.cuh"
#include "internal.h"
#include "runtime.cuh"

namespace b200 {
namespace {

constexpr int kMaxCols = B200_NOALLOC_MAX_COLS;
constexpr int kMaxCtas = 1024;
constexpr int kThreads = 256;
constexpr int kTicketDoubles = 32;  // one 256-byte line of its own in front of the partials
constexpr size_t kScratchDoubles = (size_t)kTicketDoubles + (size_t)kMaxCtas * kMaxCols;

// In host code this name is a host-side stand-in, NOT a device address: the kernel gets the address that
// cudaGetSymbolAddress returns for the current device (scratch_for_device below).
static __device__ __align__(256) double g_noalloc_scratch[kScratchDoubles];

// sum over the CTA, result valid in thread 0 (fixed tree: the same inputs give the same bits)
__device__ __forceinline__ double cta_sum(double v, double* sh) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();  // (sh may still be read from the previous use)
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  double s = 0.0;
  if (warp == 0) {
    s = lane < kThreads / 32 ? sh[lane] : 0.0;
    s = warp_sum(s);
  }
  return s;
}

// out[j] = sum_i x(i, j) y(i, j), j < ncols.  CTA b owns a contiguous range of rows for all columns.
__global__ void __launch_bounds__(kThreads)
    dot_noalloc_kernel(const double* __restrict__ x, int64_t ldx, const double* __restrict__ y, int64_t ldy,
                       int64_t n, int ncols, double* __restrict__ out, double* scratch) {
  __shared__ double sh[kThreads / 32];
  __shared__ bool last;
  unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch);
  double* part = scratch + kTicketDoubles;  // [CTA][ncols]
  const int64_t r_begin = n * blockIdx.x / gridDim.x, r_end = n * (blockIdx.x + 1) / gridDim.x;
  for (int j = 0; j < ncols; ++j) {
    const double* xj = x + (size_t)j * ldx;
    const double* yj = y + (size_t)j * ldy;
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
    int64_t i = r_begin + threadIdx.x;
    for (; i + 3 * kThreads < r_end; i += 4 * kThreads) {
      a0 = fma(xj[i], yj[i], a0);
      a1 = fma(xj[i + kThreads], yj[i + kThreads], a1);
      a2 = fma(xj[i + 2 * kThreads], yj[i + 2 * kThreads], a2);
      a3 = fma(xj[i + 3 * kThreads], yj[i + 3 * kThreads], a3);
    }
    for (; i < r_end; i += kThreads) a0 = fma(xj[i], yj[i], a0);
    const double s = cta_sum((a0 + a1) + (a2 + a3), sh);
    if (threadIdx.x == 0) part[(size_t)blockIdx.x * ncols + j] = s;
  }
  // the CTA that takes the last ticket folds the partials in CTA order
  if (threadIdx.x == 0) {
    __threadfence();
    last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  for (int j = 0; j < ncols; ++j) {
    double a = 0.0;
    for (int b = threadIdx.x; b < (int)gridDim.x; b += kThreads) a += __ldcg(&part[(size_t)b * ncols + j]);
    const double s = cta_sum(a, sh);
    if (threadIdx.x == 0) out[j] = s;
  }
  if (threadIdx.x == 0) *ticket = 0u;  // (and cleared again in front of the next launch)
}

struct DevScratch {
  double* ptr = nullptr;
  int num_sms = 0;
};
constexpr int kMaxDevices = 64;
DevScratch g_dev[kMaxDevices];
std::mutex g_dev_mu;

// set-up, once per device, with that device current: the device address of the array
int scratch_for_device(int dev, DevScratch* out) {
  std::lock_guard<std::mutex> lk(g_dev_mu);
  if (!g_dev[dev].ptr) {
    double* p = nullptr;
    B200_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
    if (!p) return fail(B200_ERR_CUDA, "b200_dot_noalloc: cudaGetSymbolAddress returned a null address");
    int sms = 0;
    B200_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    g_dev[dev].num_sms = sms > 0 ? sms : 1;
    g_dev[dev].ptr = p;
  }
  *out = g_dev[dev];
  return 0;
}

}  // namespace
}  // namespace b200

extern "C" int b200_dot_noalloc(const double* x, int64_t ldx, const double* y, int64_t ldy, int64_t n, int ncols,
                                double* out, void* stream) {
  using namespace b200;
  if (ncols < 0 || n < 0) return fail(B200_ERR_ARG, "b200_dot_noalloc: negative size");
  if (ncols > kMaxCols)
    return failf(B200_ERR_ARG, "b200_dot_noalloc: %d columns, the static scratch array holds at most %d", ncols,
                 kMaxCols);
  if (ncols == 0) return 0;
  if (!out || (n > 0 && (!x || !y))) return fail(B200_ERR_ARG, "b200_dot_noalloc: null pointer");
  if (ncols > 1 && (ldx < n || ldy < n)) return fail(B200_ERR_SHAPE, "b200_dot_noalloc: leading dimension < n");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    (void)cudaGetLastError();
    return fail(B200_ERR_NO_DEVICE, "b200_dot_noalloc: no CUDA device (there is no CPU fallback)");
  }
  int dev = 0;
  B200_CUDA(cudaGetDevice(&dev));  // the caller's current device: the one its pointers and stream belong to
  if (dev < 0 || dev >= kMaxDevices) return fail(B200_ERR_ARG, "b200_dot_noalloc: device ordinal out of range");
  DevScratch ds;
  B200_TRY(scratch_for_device(dev, &ds));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // ~8 K rows per CTA at least, one wave of 4 CTAs per SM at most
  int64_t grid = (n + 8191) / 8192;
  if (grid > 4 * (int64_t)ds.num_sms) grid = 4 * (int64_t)ds.num_sms;
  if (grid > kMaxCtas) grid = kMaxCtas;
  if (grid < 1) grid = 1;
  B200_CUDA(cudaMemsetAsync(ds.ptr, 0, kTicketDoubles * sizeof(double), st));  // the ticket must start at zero
  dot_noalloc_kernel<<<(unsigned)grid, kThreads, 0, st>>>(x, ldx, y, ldy, n, ncols, out, ds.ptr);
  B200_CUDA(cudaGetLastError());
  return 0;
}
