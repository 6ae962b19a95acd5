// MvTimesMatAddMv: V(n x q) <- alpha * A(n x p) * B(p x q) + beta * V, B a small host block.
// Second-heaviest kernel of the orthogonalisation (reference: 38 % of the config-2 solve,
// SURVEY 8a row a6).
//
// Reference semantics: Belos::EpetraMultiVec::MvTimesMatAddMv belos/epetra/src/BelosEpetraAdapter.cpp:421-435
//   -> Epetra_MultiVector::Multiply('N','N',alpha,A,B,beta)    epetra/src/Epetra_MultiVector.cpp:1936-2083
//   (beta == 0 overwrites without reading V; DGEMM semantics.)
//
// Design (HBM-bound, algorithmic bytes 8 n (p + 2q), or 8 n (p + q) when beta == 0):
//   * a thread owns 4 consecutive rows (one 256-bit load per column of A) and all q outputs
//     of those rows: 4*Q FP64 accumulators; A is streamed once with evict-first loads,
//     unrolled x8 so every thread keeps 8 independent 32-byte requests in flight.
//   * B is staged through shared memory in chunks of PCH rows, transposed to [j][q] so the
//     inner loop reads it as warp-wide broadcasts.
#include "device_utils.cuh"
#include "runtime.cuh"

namespace b200 {

namespace {

constexpr int kThreads = 128;
constexpr int kPChunk = 256;  // rows of B per shared-memory stage

template <int Q, int UNR>
__global__ void __launch_bounds__(kThreads, (Q <= 4 ? 4 : 3))
    update_kernel(double* __restrict__ V, int64_t ldv, const double* __restrict__ A, int64_t lda,
                  int p, int64_t n, const double* __restrict__ B, int ldb, double alpha,
                  double beta) {
  __shared__ double Bs[kPChunk][Q];
  const int64_t r = ((int64_t)blockIdx.x * kThreads + threadIdx.x) * 4;
  const bool active = r < n;
  const double* Ar = A + (active ? r : 0);
  double acc[4][Q];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int k = 0; k < Q; ++k) acc[i][k] = 0.0;

  for (int j0 = 0; j0 < p; j0 += kPChunk) {
    const int pc = min(kPChunk, p - j0);
    __syncthreads();
    for (int e = threadIdx.x; e < pc * Q; e += kThreads) {
      const int k = e / pc, j = e % pc;  // coalesced read of column k of B
      Bs[j][k] = B[(size_t)k * ldb + j0 + j];
    }
    __syncthreads();
    if (active) {
      const double* Aj = Ar + (size_t)j0 * lda;
      int j = 0;
      for (; j + UNR <= pc; j += UNR) {
        double4v a[UNR];
#pragma unroll
        for (int u = 0; u < UNR; ++u) a[u] = ld_stream4(Aj + (size_t)(j + u) * lda);
#pragma unroll
        for (int u = 0; u < UNR; ++u)
#pragma unroll
          for (int k = 0; k < Q; ++k) {
            const double b = Bs[j + u][k];
            acc[0][k] = fma(a[u].x, b, acc[0][k]);
            acc[1][k] = fma(a[u].y, b, acc[1][k]);
            acc[2][k] = fma(a[u].z, b, acc[2][k]);
            acc[3][k] = fma(a[u].w, b, acc[3][k]);
          }
      }
      for (; j < pc; ++j) {
        const double4v a = ld_stream4(Aj + (size_t)j * lda);
#pragma unroll
        for (int k = 0; k < Q; ++k) {
          const double b = Bs[j][k];
          acc[0][k] = fma(a.x, b, acc[0][k]);
          acc[1][k] = fma(a.y, b, acc[1][k]);
          acc[2][k] = fma(a.z, b, acc[2][k]);
          acc[3][k] = fma(a.w, b, acc[3][k]);
        }
      }
    }
  }
  if (!active) return;
#pragma unroll
  for (int k = 0; k < Q; ++k) {
    double* vp = V + (size_t)k * ldv + r;
    double4v o;
    if (beta == 0.0) {
      o.x = alpha * acc[0][k]; o.y = alpha * acc[1][k];
      o.z = alpha * acc[2][k]; o.w = alpha * acc[3][k];
    } else {
      const double4v v = ld_plain4(vp);
      o.x = fma(alpha, acc[0][k], beta * v.x);
      o.y = fma(alpha, acc[1][k], beta * v.y);
      o.z = fma(alpha, acc[2][k], beta * v.z);
      o.w = fma(alpha, acc[3][k], beta * v.w);
    }
    st_plain4(vp, o);
  }
}

template <int Q, int UNR>
int launch_update_t(Ctx* ctx, double* V, int64_t ldv, const double* A, int64_t lda, int p,
                    int64_t n, const double* dB, int ldb, double alpha, double beta) {
  const int64_t nthr = (n + 3) / 4;
  const unsigned grid = (unsigned)((nthr + kThreads - 1) / kThreads);
  update_kernel<Q, UNR><<<grid, kThreads, 0, ctx->stream>>>(V, ldv, A, lda, p, n, dB, ldb, alpha,
                                                             beta);
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace

int launch_update(Ctx* ctx, double* V, int64_t ldv, int q, const double* A, int64_t lda, int p,
                  int64_t n, const double* dB, int ldb, double alpha, double beta) {
  if (n <= 0 || q <= 0) return 0;
  ctx->t_begin(K_UPDATE, 8.0 * n * ((double)p + (beta == 0.0 ? 1.0 : 2.0) * q));
  int k0 = 0, rc = 0;
  while (k0 < q && rc == 0) {
    const int rem = q - k0;
    double* Vk = V + (size_t)k0 * ldv;
    const double* Bk = dB + (size_t)k0 * ldb;
    int step;
    if (rem >= 8) step = 8; else if (rem >= 4) step = 4; else if (rem >= 2) step = 2; else step = 1;
    switch (step) {
      case 8: rc = launch_update_t<8, 4>(ctx, Vk, ldv, A, lda, p, n, Bk, ldb, alpha, beta); break;
      case 4: rc = launch_update_t<4, 8>(ctx, Vk, ldv, A, lda, p, n, Bk, ldb, alpha, beta); break;
      case 2: rc = launch_update_t<2, 8>(ctx, Vk, ldv, A, lda, p, n, Bk, ldb, alpha, beta); break;
      default: rc = launch_update_t<1, 8>(ctx, Vk, ldv, A, lda, p, n, Bk, ldb, alpha, beta); break;
    }
    k0 += step;
  }
  ctx->t_end();
  return rc;
}

}  // namespace b200
