// Operator apply: Y = A X, FP64 CSR SpMV / SpMM over the block of right-hand sides.
//
// Reference semantics: Epetra_CrsMatrix::Multiply(false,X,Y) epetra/src/Epetra_CrsMatrix.cpp:3107-3285
//   kernels GeneralMV :3315-3413 / GeneralMM :3486-3585 (C++ loops :3352-3363, :3516-3534;
//   Fortran epetra_dcrsmv / epetra_scrsmm5, epetra/src/Epetra_dcrsmv.f:152-191,
//   Epetra_dcrsmm.f:340-500): per row, sum of val*x[col] accumulated left to right.
//
// Design (HBM-bound, algorithmic bytes 12 nnz + 4 (n+1) + 16 n k):
//   * "CSR-stream": a CTA owns 256 consecutive rows; their value / column-index segments are
//     contiguous in memory and are staged into shared memory with fully coalesced loads, then
//     each thread reduces its own row out of shared memory for all K right-hand sides
//     (K FP64 accumulators, same left-to-right order as the reference).
//   * x is gathered through L1/L2 (stencil neighbours of consecutive rows coalesce); owned
//     columns come from the caller's multivector, ghost columns (multi-rank) from the halo
//     buffer, so no column-map copy of x is ever made.
//   * interior pass: rows touching a ghost column are skipped (written later by the boundary
//     kernel once the NCCL halo exchange, which overlaps this pass, has landed).
#include "device_utils.cuh"
#include "runtime.cuh"

namespace b200 {

namespace {

constexpr int kRows = 256;     // rows per CTA == threads per CTA
constexpr int kCap = 4064;     // staged nonzeros per pass: 4064 * 12 B = 47.6 KB (+ ticket flag < 48 KB)

// FUSE: additionally emits dots[k] = sum y(:,k)^2 — the MvDot(X,X) DGKS takes of the fresh
// Krylov block right after the apply (BelosDGKSOrthoManager.hpp:1269,1395).
template <int K, bool FUSE>
__global__ void __launch_bounds__(kRows)
    spmm_stream_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ colind,
                       const double* __restrict__ values, int64_t n_rows, int32_t n_owned,
                       const double* __restrict__ x, int64_t ldx, const double* __restrict__ xg,
                       int64_t ldg, double* __restrict__ y, int64_t ldy, int skip_ghost_rows,
                       FusedDots fd) {
  __shared__ double sv[kCap];
  __shared__ int32_t sc[kCap];
  // persistent CTAs walk the 256-row tiles round-robin (see update.cu: the fused reduction's
  // fence + ticket round trip is then paid once per CTA, not once per wave of CTAs)
  const int64_t ntiles = (n_rows + kRows - 1) / kRows;
  double ss[K];
#pragma unroll
  for (int k = 0; k < K; ++k) ss[k] = 0.0;

  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t r0 = tile * kRows;
    const int64_t r = r0 + threadIdx.x;
    const int64_t rl = min(r0 + (int64_t)kRows, n_rows);
    const int32_t seg_b = rowptr[r0], seg_e = rowptr[rl];
    int32_t my_b = 0, my_e = 0;
    if (r < n_rows) {
      my_b = rowptr[r];
      my_e = rowptr[r + 1];
    }
    double acc[K];
#pragma unroll
    for (int k = 0; k < K; ++k) acc[k] = 0.0;
    bool has_ghost = false;

    for (int32_t base = seg_b; base < seg_e; base += kCap) {
      const int32_t cnt = min(kCap, seg_e - base);
      __syncthreads();
      for (int32_t i = threadIdx.x; i < cnt; i += kRows) {
        sv[i] = ld_stream1(values + base + i);
        sc[i] = __ldg(colind + base + i);
      }
      __syncthreads();
      const int32_t lo = max(my_b, base), hi = min(my_e, base + cnt);
      for (int32_t e = lo; e < hi; ++e) {
        const double v = sv[e - base];
        const int32_t c = sc[e - base];
        if (c < n_owned) {
#pragma unroll
          for (int k = 0; k < K; ++k) acc[k] = fma(v, __ldg(x + (size_t)k * ldx + c), acc[k]);
        } else {
          has_ghost = true;
          if (!skip_ghost_rows) {
#pragma unroll
            for (int k = 0; k < K; ++k)
              acc[k] = fma(v, __ldg(xg + (size_t)k * ldg + (c - n_owned)), acc[k]);
          }
        }
      }
    }
    if (r < n_rows && !(skip_ghost_rows && has_ghost)) {
#pragma unroll
      for (int k = 0; k < K; ++k) {
        y[(size_t)k * ldy + r] = acc[k];
        if (FUSE) ss[k] = fma(acc[k], acc[k], ss[k]);
      }
    }
  }
  if (!FUSE) return;
  __syncthreads();  // sv is reused as reduction scratch
  double(*red)[K] = reinterpret_cast<double(*)[K]>(sv);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    const double t = warp_sum(ss[k]);
    if (lane == 0) red[warp][k] = t;
  }
  __syncthreads();
  if (threadIdx.x < K) {
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < kRows / 32; ++w) t += red[w][threadIdx.x];
    fd.ws[(size_t)threadIdx.x * gridDim.x + blockIdx.x] = t;
  }
  if (!fold_partials_wide<kRows / 32>(fd.ws, fd.ticket, K, gridDim.x, blockIdx.x, 1, fd.out)) return;
  publish_flag(fd.flag, fd.seq);
}

// boundary rows (explicit list): one thread per row, direct loads
template <int K>
__global__ void __launch_bounds__(256)
    spmm_rows_kernel(const int32_t* __restrict__ rows, int64_t nrows,
                     const int32_t* __restrict__ rowptr, const int32_t* __restrict__ colind,
                     const double* __restrict__ values, int32_t n_owned,
                     const double* __restrict__ x, int64_t ldx, const double* __restrict__ xg,
                     int64_t ldg, double* __restrict__ y, int64_t ldy) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nrows) return;
  const int32_t r = rows[i];
  double acc[K];
#pragma unroll
  for (int k = 0; k < K; ++k) acc[k] = 0.0;
  for (int32_t e = rowptr[r]; e < rowptr[r + 1]; ++e) {
    const double v = values[e];
    const int32_t c = colind[e];
    const double* src = (c < n_owned) ? x + c : xg + (c - n_owned);
    const int64_t ld = (c < n_owned) ? ldx : ldg;
#pragma unroll
    for (int k = 0; k < K; ++k) acc[k] = fma(v, __ldg(src + (size_t)k * ld), acc[k]);
  }
#pragma unroll
  for (int k = 0; k < K; ++k) y[(size_t)k * ldy + r] = acc[k];
}

// halo pack: buf[j*cnt + i] = x[idx[i] + j*ldx]   (Epetra PackAndPrepare,
// epetra/src/Epetra_MultiVector.cpp:724-813, but RHS-major so each (peer, rhs) slice is contiguous)
__global__ void __launch_bounds__(256)
    pack_kernel(const int32_t* __restrict__ idx, int64_t cnt, const double* __restrict__ x,
                int64_t ldx, double* __restrict__ buf) {
  const int j = blockIdx.y;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < cnt;
       i += (int64_t)gridDim.x * blockDim.x)
    buf[(size_t)j * cnt + i] = __ldg(x + (size_t)j * ldx + idx[i]);
}

template <int K>
int spmm_t(Ctx* ctx, const Op* op, const int32_t* rows, int64_t nrows, const double* x,
           int64_t ldx, const double* xg, int64_t ldg, double* y, int64_t ldy, int skip,
           const FusedDots& fd) {
  if (rows == nullptr) {
    const int64_t ntiles = (op->n_local + kRows - 1) / kRows;
    if (ntiles == 0) return 0;
    static int per_sm = 0;  // resident CTAs per SM of this instantiation
    if (per_sm == 0) {
      B200_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, spmm_stream_kernel<K, true>,
                                                              kRows, 0));
      per_sm = std::max(per_sm, 1);
    }
    const unsigned grid = (unsigned)std::min<int64_t>(ntiles, (int64_t)ctx->num_sms * per_sm);
    if (fd.out)
      spmm_stream_kernel<K, true><<<grid, kRows, 0, ctx->stream>>>(
          op->rowptr, op->colind, op->values, op->n_local, (int32_t)op->n_local, x, ldx, xg, ldg,
          y, ldy, skip, fd);
    else
      spmm_stream_kernel<K, false><<<grid, kRows, 0, ctx->stream>>>(
          op->rowptr, op->colind, op->values, op->n_local, (int32_t)op->n_local, x, ldx, xg, ldg,
          y, ldy, skip, fd);
  } else {
    if (nrows == 0) return 0;
    const unsigned grid = (unsigned)((nrows + 255) / 256);
    spmm_rows_kernel<K><<<grid, 256, 0, ctx->stream>>>(rows, nrows, op->rowptr, op->colind,
                                                        op->values, (int32_t)op->n_local, x, ldx,
                                                        xg, ldg, y, ldy);
  }
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace

// rows == nullptr: stream kernel over all local rows (skipping ghost rows when the operator has
// ghosts and the caller asked for the interior pass via nrows == -1); otherwise the row list.
int launch_spmm(Ctx* ctx, const Op* op, const int32_t* rows, int64_t nrows, const double* x,
                int64_t ldx, double* y, int64_t ldy, int k, const FusedDots* fused) {
  const int skip = (rows == nullptr && nrows == -1) ? 1 : 0;
  const double* xg = op->recv_buf;
  const int64_t ldg = op->n_ghost;  // ghost block is stored densely per right-hand side
  int k0 = 0;
  while (k0 < k) {
    const int rem = k - k0;
    int step;
    if (rem >= 8) step = 8; else if (rem >= 4) step = 4; else if (rem >= 2) step = 2; else step = 1;
    const double* xk = x + (size_t)k0 * ldx;
    const double* xgk = xg ? xg + (size_t)k0 * ldg : nullptr;
    double* yk = y + (size_t)k0 * ldy;
    int rc;
    FusedDots fd;
    if (fused && fused->out && rows == nullptr && !skip) {
      fd = *fused;
      fd.out = fused->out + k0;
      if (k0 + step < k) fd.flag = nullptr;
    }
    switch (step) {
      case 8: rc = spmm_t<8>(ctx, op, rows, nrows, xk, ldx, xgk, ldg, yk, ldy, skip, fd); break;
      case 4: rc = spmm_t<4>(ctx, op, rows, nrows, xk, ldx, xgk, ldg, yk, ldy, skip, fd); break;
      case 2: rc = spmm_t<2>(ctx, op, rows, nrows, xk, ldx, xgk, ldg, yk, ldy, skip, fd); break;
      default: rc = spmm_t<1>(ctx, op, rows, nrows, xk, ldx, xgk, ldg, yk, ldy, skip, fd); break;
    }
    if (rc) return rc;
    k0 += step;
  }
  return 0;
}

int launch_pack(Ctx* ctx, cudaStream_t s, const int32_t* idx, int64_t cnt, const double* x,
                int64_t ldx, int k, double* buf) {
  if (cnt <= 0 || k <= 0) return 0;
  const unsigned bx = (unsigned)std::min<int64_t>((cnt + 255) / 256, kNumSM * 4);
  pack_kernel<<<dim3(bx, (unsigned)k), 256, 0, s>>>(idx, cnt, x, ldx, buf);
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace b200
