// MvTransMv: C(p x q) = alpha * A^T V for a tall-skinny Krylov basis A (n x p) and a block
// V (n x q), both FP64 column-major.  This is the kernel that dominates Block GMRES
// orthogonalisation (reference: 46 % of the config-2 solve).
//
// Reference semantics: Belos::EpetraMultiVec::MvTransMv  belos/epetra/src/BelosEpetraAdapter.cpp:480-495
//   -> Epetra_MultiVector::Multiply('T','N',alpha,A,V,0)   epetra/src/Epetra_MultiVector.cpp:1936-2083
//   -> local DGEMM + Reduce() (SumAll)                     :2400-2433
//
// Design (HBM-bound, algorithmic bytes 8 n (p + q)):
//   * A is streamed from HBM exactly once with 256-bit evict-first loads; V (n x q, <= a few
//     tens of MB) is kept L2/L1 resident with evict-last loads and re-read by every column block.
//   * grid = (row ranges) x (column blocks); inside a CTA the 8 warps are arranged as
//     WC column groups x WR row groups; a warp owns CPW columns of A and all Q columns of V,
//     a lane owns 4 consecutive rows per step, so it keeps CPW*Q FP64 accumulators for the
//     whole row range (no per-tile reduction).
//   * per-CTA partials go to a workspace laid out [entry][part]; the CTA that arrives last
//     folds them in a fixed order (deterministic) and publishes to the host mailbox.
#include "device_utils.cuh"
#include "runtime.cuh"

namespace b200 {

namespace {

constexpr int kThreads = 256;
constexpr int kWarps = 8;
constexpr int kRowsPerWarpStep = 128;  // 32 lanes x 4 rows

template <int Q, int CPW>
__global__ void __launch_bounds__(kThreads, 2)
    gram_kernel(const double* __restrict__ A, int64_t lda, int p, const double* __restrict__ V,
                int64_t ldv, int64_t n, int64_t rows_per_cta, int wc_log2, double alpha,
                double* __restrict__ ws, int nparts_pad, unsigned int* ticket,
                double* __restrict__ out, int ldc, uint64_t* flag, uint64_t seq) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int WC = 1 << wc_log2, WR = kWarps >> wc_log2;
  const int wc = warp & (WC - 1), wr = warp >> wc_log2;
  const int col0 = (blockIdx.y * WC + wc) * CPW;
  const int64_t r_begin = (int64_t)blockIdx.x * rows_per_cta;
  const int64_t r_end = min(n, r_begin + rows_per_cta);

  double acc[CPW][Q];
#pragma unroll
  for (int c = 0; c < CPW; ++c)
#pragma unroll
    for (int k = 0; k < Q; ++k) acc[c][k] = 0.0;

  const double* Ap[CPW];
#pragma unroll
  for (int c = 0; c < CPW; ++c) Ap[c] = A + (size_t)min(col0 + c, p - 1) * lda;

  if (col0 < p) {
    for (int64_t r = r_begin + (int64_t)wr * kRowsPerWarpStep + lane * 4; r < r_end;
         r += (int64_t)kRowsPerWarpStep * WR) {
      double4v a[CPW], v[Q];
#pragma unroll
      for (int c = 0; c < CPW; ++c) a[c] = ld_stream4(Ap[c] + r);
#pragma unroll
      for (int k = 0; k < Q; ++k) v[k] = ld_keep4(V + (size_t)k * ldv + r);
      if (r + 3 >= n) {  // last, partial 4-row vector: padding rows must not contribute
#pragma unroll
        for (int c = 0; c < CPW; ++c) {
          if (r + 1 >= n) a[c].y = 0.0;
          if (r + 2 >= n) a[c].z = 0.0;
          if (r + 3 >= n) a[c].w = 0.0;
        }
#pragma unroll
        for (int k = 0; k < Q; ++k) {
          if (r + 1 >= n) v[k].y = 0.0;
          if (r + 2 >= n) v[k].z = 0.0;
          if (r + 3 >= n) v[k].w = 0.0;
        }
      }
#pragma unroll
      for (int c = 0; c < CPW; ++c)
#pragma unroll
        for (int k = 0; k < Q; ++k) {
          acc[c][k] = fma(a[c].x, v[k].x, acc[c][k]);
          acc[c][k] = fma(a[c].y, v[k].y, acc[c][k]);
          acc[c][k] = fma(a[c].z, v[k].z, acc[c][k]);
          acc[c][k] = fma(a[c].w, v[k].w, acc[c][k]);
        }
    }
  }

  // CTA reduction: lanes -> warp, then the WR row-group warps of each column group
  __shared__ double sm[kWarps][CPW * Q];
#pragma unroll
  for (int c = 0; c < CPW; ++c)
#pragma unroll
    for (int k = 0; k < Q; ++k) {
      const double s = warp_sum(acc[c][k]);
      if (lane == 0) sm[warp][c * Q + k] = s;
    }
  __syncthreads();
  const int nparts = gridDim.x;
  for (int e = threadIdx.x; e < WC * CPW * Q; e += kThreads) {
    const int g = e / (CPW * Q), ck = e % (CPW * Q);
    const int c = ck / Q, k = ck % Q;
    const int col = (blockIdx.y * WC + g) * CPW + c;
    if (col < p) {
      double s = 0.0;
      for (int w = 0; w < WR; ++w) s += sm[(w << wc_log2) + g][ck];
      ws[((size_t)k * p + col) * nparts_pad + blockIdx.x] = s;
    }
  }

  if (!last_cta_arrives(ticket, gridDim.x * gridDim.y)) return;
  // deterministic fold over the row-range partials: one warp per output entry
  const int nent = p * Q;
  for (int e = warp; e < nent; e += kWarps) {
    const volatile double* src = ws + (size_t)e * nparts_pad;
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    int i = lane;
    for (; i + 96 < nparts; i += 128) {
      s0 += src[i]; s1 += src[i + 32]; s2 += src[i + 64]; s3 += src[i + 96];
    }
    for (; i < nparts; i += 32) s0 += src[i];
    const double s = warp_sum((s0 + s1) + (s2 + s3));
    if (lane == 0) {
      const int k = e / p, col = e % p;
      out[(size_t)k * ldc + col] = alpha * s;
      fence_results();
    }
  }
  publish_flag(flag, seq);
}

template <int Q, int CPW>
int launch_gram_t(Ctx* ctx, const double* A, int64_t lda, int p, const double* V, int64_t ldv,
                  int64_t n, double alpha, double* out, int ldc, uint64_t* flag) {
  // warp arrangement: as many column groups as the block needs (power of two <= 8)
  int wc_log2 = 0;
  while ((1 << wc_log2) < kWarps && (1 << wc_log2) * CPW < p) ++wc_log2;
  const int WC = 1 << wc_log2, WR = kWarps >> wc_log2;
  const int tc = WC * CPW;
  const int ncb = (p + tc - 1) / tc;
  const int64_t step = (int64_t)kRowsPerWarpStep * WR;
  int target = std::max(1, (2 * kNumSM) / ncb);
  int64_t rows_per_cta = (n + target - 1) / target;
  rows_per_cta = std::max<int64_t>(step, (rows_per_cta + step - 1) / step * step);
  int nparts = (int)std::max<int64_t>(1, (n + rows_per_cta - 1) / rows_per_cta);
  const int nparts_pad = (nparts + 31) & ~31;
  if ((size_t)p * Q * nparts_pad * sizeof(double) > kWorkspaceBytes)
    return fail(B200_ERR_ARG, "MvTransMv: partial-sum workspace too small");
  dim3 grid((unsigned)nparts, (unsigned)ncb, 1);
  gram_kernel<Q, CPW><<<grid, kThreads, 0, ctx->stream>>>(A, lda, p, V, ldv, n, rows_per_cta,
                                                           wc_log2, alpha, ctx->ws, nparts_pad,
                                                           ctx->ticket, out, ldc, flag, ctx->seq);
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace

int launch_gram(Ctx* ctx, const double* A, int64_t lda, int p, const double* V, int64_t ldv, int q,
                int64_t n, double alpha, int out_off, int ldc, bool last) {
  if (p <= 0 || q <= 0) return 0;
  double* out = ctx->result_ptr() + out_off;
  ctx->t_begin(K_GRAM, 8.0 * n * ((double)p + q));
  // split q greedily into register-resident chunks of 8 / 4 / 2 / 1 columns of V
  int k0 = 0;
  int rc = 0;
  while (k0 < q && rc == 0) {
    const int rem = q - k0;
    const double* Vk = V + (size_t)k0 * ldv;
    double* outk = out + (size_t)k0 * ldc;
    int step;
    if (rem >= 8) step = 8; else if (rem >= 4) step = 4; else if (rem >= 2) step = 2; else step = 1;
    uint64_t* flag = (last && k0 + step >= q) ? ctx->result_flag() : nullptr;
    switch (step) {
      case 8: rc = launch_gram_t<8, 2>(ctx, A, lda, p, Vk, ldv, n, alpha, outk, ldc, flag); break;
      case 4: rc = launch_gram_t<4, 4>(ctx, A, lda, p, Vk, ldv, n, alpha, outk, ldc, flag); break;
      case 2: rc = launch_gram_t<2, 4>(ctx, A, lda, p, Vk, ldv, n, alpha, outk, ldc, flag); break;
      default: rc = launch_gram_t<1, 4>(ctx, A, lda, p, Vk, ldv, n, alpha, outk, ldc, flag); break;
    }
    k0 += step;
  }
  ctx->t_end();
  return rc;
}

}  // namespace b200
