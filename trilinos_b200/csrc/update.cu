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

// FUSE: the kernel additionally emits, per output column k, dots[2k] = sum new(:,k)^2 and
// dots[2k+1] = sum new(:,k)*old(:,k) (old = V before the update; only when beta != 0) — the two
// inner products DGKS asks for right after the update (BelosDGKSOrthoManager.hpp:1091,1128,1480),
// so those MvDot calls need neither a launch nor another pass over the column.
template <int Q, int UNR, bool FUSE>
__global__ void __launch_bounds__(kThreads, (Q <= 4 ? 4 : 3))
    update_kernel(double* __restrict__ V, int64_t ldv, const double* __restrict__ A, int64_t lda,
                  int p, int64_t n, const double* __restrict__ B, int ldb, double alpha,
                  double beta, FusedDots fd) {
  __shared__ double Bs[kPChunk][Q];
  // Persistent CTAs (grid = resident capacity) walk the 512-row tiles round-robin: the fused
  // reduction's fence + ticket round trip (~2 us during which the CTA streams nothing) is paid
  // once per CTA instead of once per wave of CTAs.
  const int64_t ntiles = (n + 4 * kThreads - 1) / (4 * kThreads);
  const bool single = p <= kPChunk;  // B fits one stage: load it once for all tiles
  double dself[Q], dold[Q];
#pragma unroll
  for (int k = 0; k < Q; ++k) {
    dself[k] = 0.0;
    dold[k] = 0.0;
  }
  auto stage = [&](int j0, int pc) {
    for (int e = threadIdx.x; e < pc * Q; e += kThreads) {
      const int k = e / pc, j = e % pc;  // coalesced read of column k of B
      Bs[j][k] = B[(size_t)k * ldb + j0 + j];
    }
  };
  if (single) {
    stage(0, p);
    __syncthreads();
  }
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t r = (tile * kThreads + threadIdx.x) * 4;
    const bool active = r < n;
    const double* Ar = A + (active ? r : 0);
    double acc[4][Q];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int k = 0; k < Q; ++k) acc[i][k] = 0.0;

    for (int j0 = 0; j0 < p; j0 += kPChunk) {
      const int pc = min(kPChunk, p - j0);
      if (!single) {
        __syncthreads();
        stage(j0, pc);
        __syncthreads();
      }
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
    if (active) {
#pragma unroll
      for (int k = 0; k < Q; ++k) {
        double* vp = V + (size_t)k * ldv + r;
        double4v o, v{0.0, 0.0, 0.0, 0.0};
        if (beta == 0.0) {
          o.x = alpha * acc[0][k]; o.y = alpha * acc[1][k];
          o.z = alpha * acc[2][k]; o.w = alpha * acc[3][k];
        } else {
          v = ld_plain4(vp);
          o.x = fma(alpha, acc[0][k], beta * v.x);
          o.y = fma(alpha, acc[1][k], beta * v.y);
          o.z = fma(alpha, acc[2][k], beta * v.z);
          o.w = fma(alpha, acc[3][k], beta * v.w);
        }
        st_plain4(vp, o);
        if (FUSE) {
          // rows >= n of the last 4-vector are padding and must not enter the sums
          const bool m1 = r + 1 < n, m2 = r + 2 < n, m3 = r + 3 < n;
          double s0 = o.x * o.x, s1 = m1 ? o.y * o.y : 0.0;
          s0 = m2 ? fma(o.z, o.z, s0) : s0;
          s1 = m3 ? fma(o.w, o.w, s1) : s1;
          dself[k] += s0 + s1;
          double t0 = o.x * v.x, t1 = m1 ? o.y * v.y : 0.0;
          t0 = m2 ? fma(o.z, v.z, t0) : t0;
          t1 = m3 ? fma(o.w, v.w, t1) : t1;
          dold[k] += t0 + t1;
        }
      }
    }
  }
  if (!FUSE) return;
  // CTA partials -> workspace [entry][cta]; two-level deterministic fold, last CTA publishes
  __shared__ double red[kThreads / 32][2 * Q];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < Q; ++k) {
    const double a = warp_sum(dself[k]), b = warp_sum(dold[k]);
    if (lane == 0) {
      red[warp][2 * k] = a;
      red[warp][2 * k + 1] = b;
    }
  }
  __syncthreads();
  if (threadIdx.x < 2 * Q) {
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) t += red[w][threadIdx.x];
    fd.ws[(size_t)threadIdx.x * gridDim.x + blockIdx.x] = t;
  }
  if (!fold_partials_wide<kThreads / 32>(fd.ws, fd.ticket, 2 * Q, gridDim.x, blockIdx.x, 1, fd.out))
    return;
  publish_flag(fd.flag, fd.seq);
}

template <int Q, int UNR>
int launch_update_t(Ctx* ctx, double* V, int64_t ldv, const double* A, int64_t lda, int p,
                    int64_t n, const double* dB, int ldb, double alpha, double beta,
                    const FusedDots& fd) {
  const int64_t ntiles = (n + 4 * kThreads - 1) / (4 * kThreads);
  static int per_sm = 0;  // resident CTAs per SM of this instantiation
  if (per_sm == 0) {
    B200_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, update_kernel<Q, UNR, true>,
                                                            kThreads, 0));
    per_sm = std::max(per_sm, 1);
  }
  const unsigned grid = (unsigned)std::min<int64_t>(ntiles, (int64_t)ctx->num_sms * per_sm);
  if (fd.out)
    update_kernel<Q, UNR, true><<<grid, kThreads, 0, ctx->stream>>>(V, ldv, A, lda, p, n, dB, ldb,
                                                                     alpha, beta, fd);
  else
    update_kernel<Q, UNR, false><<<grid, kThreads, 0, ctx->stream>>>(V, ldv, A, lda, p, n, dB,
                                                                      ldb, alpha, beta, fd);
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace

int launch_update(Ctx* ctx, double* V, int64_t ldv, int q, const double* A, int64_t lda, int p,
                  int64_t n, const double* dB, int ldb, double alpha, double beta,
                  const FusedDots* fused) {
  if (n <= 0 || q <= 0) return 0;
  ctx->t_begin(K_UPDATE, 8.0 * n * ((double)p + (beta == 0.0 ? 1.0 : 2.0) * q));
  int k0 = 0, rc = 0;
  while (k0 < q && rc == 0) {
    const int rem = q - k0;
    double* Vk = V + (size_t)k0 * ldv;
    const double* Bk = dB + (size_t)k0 * ldb;
    int step;
    if (rem >= 8) step = 8; else if (rem >= 4) step = 4; else if (rem >= 2) step = 2; else step = 1;
    FusedDots fd;
    if (fused && fused->out) {
      fd = *fused;
      fd.out = fused->out + 2 * k0;
      if (k0 + step < q) fd.flag = nullptr;  // only the last launch releases the sequence number
    }
    switch (step) {
      case 8: rc = launch_update_t<8, 4>(ctx, Vk, ldv, A, lda, p, n, Bk, ldb, alpha, beta, fd); break;
      case 4: rc = launch_update_t<4, 8>(ctx, Vk, ldv, A, lda, p, n, Bk, ldb, alpha, beta, fd); break;
      case 2: rc = launch_update_t<2, 8>(ctx, Vk, ldv, A, lda, p, n, Bk, ldb, alpha, beta, fd); break;
      default: rc = launch_update_t<1, 8>(ctx, Vk, ldv, A, lda, p, n, Bk, ldb, alpha, beta, fd); break;
    }
    k0 += step;
  }
  ctx->t_end();
  return rc;
}

}  // namespace b200
