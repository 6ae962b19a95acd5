// Streaming element-wise multivector kernels (MvAddMv, MvScale, MvInit, MvRandom, copies).
// All are HBM/L2-bandwidth bound: 256-bit loads/stores, one column per blockIdx.y,
// grid-stride over 4-row vectors.  Columns are 256-byte aligned (ld % 32 == 0), so touching
// the padding rows [n, round4(n)) of a column is safe and never observable.
//
// Reference semantics: Epetra_MultiVector::Update  epetra/src/Epetra_MultiVector.cpp:1276-1500
//                      Epetra_MultiVector::Scale   :1229-1248, PutScalar :595-608,
//                      Random :490-515 with Epetra_Util::RandomDouble epetra/src/Epetra_Util.cpp:71-99
#include "device_utils.cuh"
#include "runtime.cuh"

namespace b200 {

namespace {

constexpr int kThreads = 256;

inline dim3 ew_grid(int64_t n, int k) {
  int64_t vecs = (n + 3) / 4;
  int64_t bx = (vecs + kThreads - 1) / kThreads;
  // enough CTAs for ~4 waves of 148 SMs x 8 resident CTAs, but no more than needed
  int64_t cap = std::max<int64_t>(1, (int64_t)kNumSM * 8 * 4 / std::max(1, k));
  if (bx > cap) bx = cap;
  if (bx < 1) bx = 1;
  return dim3((unsigned)bx, (unsigned)k, 1);
}

__global__ void __launch_bounds__(kThreads) fill_kernel(double* __restrict__ y, int64_t ld,
                                                         int64_t nvec, double a) {
  double* yc = y + (size_t)blockIdx.y * ld;
  const double4v v{a, a, a, a};
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nvec;
       i += (int64_t)gridDim.x * blockDim.x)
    st_plain4(yc + 4 * i, v);
}

// MODE 0: y = alpha*A ; MODE 1: y = alpha*A + beta*B.  A/B may alias y exactly.
template <int MODE>
__global__ void __launch_bounds__(kThreads)
    axpby_kernel(double* y, int64_t ldy, const double* a, int64_t lda, const double* b,
                 int64_t ldb, int64_t nvec, double alpha, double beta) {
  double* yc = y + (size_t)blockIdx.y * ldy;
  const double* ac = a + (size_t)blockIdx.y * lda;
  const double* bc = (MODE == 1) ? b + (size_t)blockIdx.y * ldb : nullptr;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nvec;
       i += (int64_t)gridDim.x * blockDim.x) {
    double4v va = ld_plain4(ac + 4 * i);
    double4v r;
    if (MODE == 0) {
      r.x = alpha * va.x; r.y = alpha * va.y; r.z = alpha * va.z; r.w = alpha * va.w;
    } else {
      double4v vb = ld_plain4(bc + 4 * i);
      r.x = alpha * va.x + beta * vb.x;
      r.y = alpha * va.y + beta * vb.y;
      r.z = alpha * va.z + beta * vb.z;
      r.w = alpha * va.w + beta * vb.w;
    }
    st_plain4(yc + 4 * i, r);
  }
}

// Element-wise product with a (broadcast) vector:  y(:,j) = sthis * y(:,j) + sab * d(:) .* b(:,j)
// (Epetra_MultiVector::Multiply(ScalarAB, A, B, ScalarThis), epetra/src/Epetra_MultiVector.cpp,
// A a single vector).  READ_Y = false never reads y (ScalarThis == 0 branch).  No FMA contraction:
// the products and the sum round as the reference's scalar loop does.
template <bool READ_Y>
__global__ void __launch_bounds__(kThreads)
    ewise_mul_kernel(double* y, int64_t ldy, const double* __restrict__ d, const double* b,
                     int64_t ldb, int64_t nvec, double sab, double sthis) {
  double* yc = y + (size_t)blockIdx.y * ldy;
  const double* bc = b + (size_t)blockIdx.y * ldb;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nvec;
       i += (int64_t)gridDim.x * blockDim.x) {
    const double4v vd = ld_plain4(d + 4 * i), vb = ld_plain4(bc + 4 * i);
    double4v r;
    if (READ_Y) {
      const double4v vy = ld_plain4(yc + 4 * i);
      r.x = __dadd_rn(__dmul_rn(sthis, vy.x), __dmul_rn(__dmul_rn(sab, vd.x), vb.x));
      r.y = __dadd_rn(__dmul_rn(sthis, vy.y), __dmul_rn(__dmul_rn(sab, vd.y), vb.y));
      r.z = __dadd_rn(__dmul_rn(sthis, vy.z), __dmul_rn(__dmul_rn(sab, vd.z), vb.z));
      r.w = __dadd_rn(__dmul_rn(sthis, vy.w), __dmul_rn(__dmul_rn(sab, vd.w), vb.w));
    } else if (sab == 1.0) {
      r.x = __dmul_rn(vd.x, vb.x); r.y = __dmul_rn(vd.y, vb.y);
      r.z = __dmul_rn(vd.z, vb.z); r.w = __dmul_rn(vd.w, vb.w);
    } else {
      r.x = __dmul_rn(__dmul_rn(sab, vd.x), vb.x); r.y = __dmul_rn(__dmul_rn(sab, vd.y), vb.y);
      r.z = __dmul_rn(__dmul_rn(sab, vd.z), vb.z); r.w = __dmul_rn(__dmul_rn(sab, vd.w), vb.w);
    }
    st_plain4(yc + 4 * i, r);
  }
}

// Chebyshev smoother steps, arithmetic as ifpack/src/Ifpack_Chebyshev.cpp (point mode, zero
// starting solution), one fused pass each instead of the reference's Scale / loop / Update:
//   first (:613-632):  w = invd * x / theta  (single vector: (invd*x)*oot; block: x*(invd*oot)); y = w
//   step  (:641-715):  w = d1*w + (x - v) * (d2*invd);  y = y + w
template <bool SINGLE>
__global__ void __launch_bounds__(kThreads)
    cheby_first_kernel(double* w, int64_t ldw, double* y, int64_t ldy,
                       const double* __restrict__ invd, const double* x, int64_t ldx, int64_t nvec,
                       double oot) {
  double* wc = w + (size_t)blockIdx.y * ldw;
  double* yc = y + (size_t)blockIdx.y * ldy;
  const double* xc = x + (size_t)blockIdx.y * ldx;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nvec;
       i += (int64_t)gridDim.x * blockDim.x) {
    const double4v vd = ld_plain4(invd + 4 * i), vx = ld_plain4(xc + 4 * i);
    double4v r;
    if (SINGLE) {
      r.x = __dmul_rn(__dmul_rn(vd.x, vx.x), oot); r.y = __dmul_rn(__dmul_rn(vd.y, vx.y), oot);
      r.z = __dmul_rn(__dmul_rn(vd.z, vx.z), oot); r.w = __dmul_rn(__dmul_rn(vd.w, vx.w), oot);
    } else {
      r.x = __dmul_rn(vx.x, __dmul_rn(vd.x, oot)); r.y = __dmul_rn(vx.y, __dmul_rn(vd.y, oot));
      r.z = __dmul_rn(vx.z, __dmul_rn(vd.z, oot)); r.w = __dmul_rn(vx.w, __dmul_rn(vd.w, oot));
    }
    st_plain4(wc + 4 * i, r);
    st_plain4(yc + 4 * i, r);
  }
}

__global__ void __launch_bounds__(kThreads)
    cheby_step_kernel(double* w, int64_t ldw, double* y, int64_t ldy,
                      const double* __restrict__ invd, const double* x, int64_t ldx,
                      const double* v, int64_t ldv, int64_t nvec, double d1, double d2) {
  double* wc = w + (size_t)blockIdx.y * ldw;
  double* yc = y + (size_t)blockIdx.y * ldy;
  const double* xc = x + (size_t)blockIdx.y * ldx;
  const double* vc = v + (size_t)blockIdx.y * ldv;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nvec;
       i += (int64_t)gridDim.x * blockDim.x) {
    const double4v vd = ld_plain4(invd + 4 * i), vx = ld_plain4(xc + 4 * i);
    const double4v vv = ld_plain4(vc + 4 * i), vw = ld_plain4(wc + 4 * i), vy = ld_plain4(yc + 4 * i);
    double4v nw, ny;
    nw.x = __dadd_rn(__dmul_rn(d1, vw.x), __dmul_rn(__dsub_rn(vx.x, vv.x), __dmul_rn(d2, vd.x)));
    nw.y = __dadd_rn(__dmul_rn(d1, vw.y), __dmul_rn(__dsub_rn(vx.y, vv.y), __dmul_rn(d2, vd.y)));
    nw.z = __dadd_rn(__dmul_rn(d1, vw.z), __dmul_rn(__dsub_rn(vx.z, vv.z), __dmul_rn(d2, vd.z)));
    nw.w = __dadd_rn(__dmul_rn(d1, vw.w), __dmul_rn(__dsub_rn(vx.w, vv.w), __dmul_rn(d2, vd.w)));
    ny.x = __dadd_rn(nw.x, vy.x); ny.y = __dadd_rn(nw.y, vy.y);
    ny.z = __dadd_rn(nw.z, vy.z); ny.w = __dadd_rn(nw.w, vy.w);
    st_plain4(wc + 4 * i, nw);
    st_plain4(yc + 4 * i, ny);
  }
}

// diag[r] = A(r, r) (0 if the row stores no diagonal entry): Epetra_CrsMatrix::ExtractDiagonalCopy;
// with min_diag >= 0 the entry is replaced by its guarded reciprocal as Ifpack_Chebyshev::Compute
// does (:345-351: |d| < min_diag -> min_diag, else 1/d)
__global__ void __launch_bounds__(kThreads)
    diag_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ colind,
                const double* __restrict__ values, int64_t n, double* __restrict__ out,
                int invert, double min_diag) {
  const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (r >= n) return;
  double d = 0.0;
  for (int32_t e = rowptr[r]; e < rowptr[r + 1]; ++e)
    if (colind[e] == (int32_t)r) {
      d = values[e];
      break;
    }
  if (invert) d = (fabs(d) < min_diag) ? min_diag : __ddiv_rn(1.0, d);
  out[r] = d;
}

struct ColScales { double a[32]; };

__global__ void __launch_bounds__(kThreads)
    scale_cols_kernel(double* y, int64_t ld, int64_t nvec, ColScales s) {
  double* yc = y + (size_t)blockIdx.y * ld;
  const double a = s.a[blockIdx.y];
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nvec;
       i += (int64_t)gridDim.x * blockDim.x) {
    double4v v = ld_plain4(yc + 4 * i);
    v.x *= a; v.y *= a; v.z *= a; v.w *= a;
    st_plain4(yc + 4 * i, v);
  }
}

// Park-Miller minimal standard generator: s <- 16807*s mod (2^31-1); value = 2*(s/m) - 1.
// Element (global row g, column c) takes the (seq_start + g + c*n_global + 1)-th state after
// `seed`, so the stream is independent of the row partition and of the launch geometry.
__device__ __forceinline__ unsigned int pm_mul(unsigned int a, unsigned int b) {
  return (unsigned int)(((unsigned long long)a * b) % 2147483647ull);
}
__device__ unsigned int pm_pow(unsigned long long e) {
  unsigned int r = 1, base = 16807u;
  while (e) {
    if (e & 1ull) r = pm_mul(r, base);
    base = pm_mul(base, base);
    e >>= 1;
  }
  return r;
}
constexpr int kRandChunk = 32;
__global__ void __launch_bounds__(kThreads)
    random_kernel(double* y, int64_t ld, int64_t n, int64_t row_offset, int64_t n_global,
                  int64_t seq_start, unsigned int seed) {
  double* yc = y + (size_t)blockIdx.y * ld;
  const int64_t nchunk = (n + kRandChunk - 1) / kRandChunk;
  for (int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; c < nchunk;
       c += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r0 = c * kRandChunk;
    const unsigned long long idx =
        (unsigned long long)seq_start + (unsigned long long)(row_offset + r0) +
        (unsigned long long)blockIdx.y * (unsigned long long)n_global;
    unsigned int s = pm_mul(seed % 2147483647u, pm_pow(idx));  // state before element r0
    const int64_t r1 = min(n, r0 + (int64_t)kRandChunk);
    for (int64_t r = r0; r < r1; ++r) {
      s = pm_mul(s, 16807u);
      yc[r] = 2.0 * ((double)s / 2147483647.0) - 1.0;
    }
  }
}

// bench hygiene: sweep a > L2-sized buffer with loads, leaving L2 full of CLEAN lines (a memset
// would leave ~126 MB of dirty lines whose write-back steals HBM bandwidth from the timed kernel)
__global__ void __launch_bounds__(kThreads) l2_sweep_kernel(const double* __restrict__ x,
                                                            int64_t nvec, double* sink) {
  double acc = 0.0;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nvec;
       i += (int64_t)gridDim.x * blockDim.x) {
    const double4v v = ld_plain4(x + 4 * i);
    acc += v.x + v.y + v.z + v.w;
  }
  if (acc == 123.456) *sink = acc;  // never true for a zero-filled buffer; keeps the loads alive
}

}  // namespace

int launch_l2_sweep(Ctx* ctx, const double* buf, size_t bytes, double* sink) {
  l2_sweep_kernel<<<kNumSM * 8, kThreads, 0, ctx->stream>>>(buf, (int64_t)(bytes / 32), sink);
  B200_CUDA(cudaGetLastError());
  return 0;
}

int launch_ewise_mul(Ctx* ctx, double* y, int64_t ldy, const double* d, const double* b, int64_t ldb,
                     int64_t n, int k, double sab, double sthis) {
  if (n <= 0 || k <= 0) return 0;
  ctx->t_begin(K_AXPBY, 8.0 * n * ((sthis == 0.0 ? 2.0 : 3.0) * k + 1));
  const dim3 g = ew_grid(n, k);
  const int64_t nvec = (n + 3) / 4;
  if (sthis == 0.0)
    ewise_mul_kernel<false><<<g, kThreads, 0, ctx->stream>>>(y, ldy, d, b, ldb, nvec, sab, sthis);
  else
    ewise_mul_kernel<true><<<g, kThreads, 0, ctx->stream>>>(y, ldy, d, b, ldb, nvec, sab, sthis);
  ctx->t_end();
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

int launch_cheby_first(Ctx* ctx, double* w, int64_t ldw, double* y, int64_t ldy, const double* invd,
                       const double* x, int64_t ldx, int64_t n, int k, double oot, bool single) {
  if (n <= 0 || k <= 0) return 0;
  ctx->t_begin(K_AXPBY, 8.0 * n * (3.0 * k + 1));
  const dim3 g = ew_grid(n, k);
  const int64_t nvec = (n + 3) / 4;
  if (single)
    cheby_first_kernel<true><<<g, kThreads, 0, ctx->stream>>>(w, ldw, y, ldy, invd, x, ldx, nvec, oot);
  else
    cheby_first_kernel<false><<<g, kThreads, 0, ctx->stream>>>(w, ldw, y, ldy, invd, x, ldx, nvec, oot);
  ctx->t_end();
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

int launch_cheby_step(Ctx* ctx, double* w, int64_t ldw, double* y, int64_t ldy, const double* invd,
                      const double* x, int64_t ldx, const double* v, int64_t ldv, int64_t n, int k,
                      double d1, double d2) {
  if (n <= 0 || k <= 0) return 0;
  ctx->t_begin(K_AXPBY, 8.0 * n * (6.0 * k + 1));
  const dim3 g = ew_grid(n, k);
  cheby_step_kernel<<<g, kThreads, 0, ctx->stream>>>(w, ldw, y, ldy, invd, x, ldx, v, ldv,
                                                     (n + 3) / 4, d1, d2);
  ctx->t_end();
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

int launch_diag(Ctx* ctx, const Op* op, double* out, bool invert, double min_diag) {
  if (op->n_local <= 0) return 0;
  const unsigned g = (unsigned)((op->n_local + kThreads - 1) / kThreads);
  diag_kernel<<<g, kThreads, 0, ctx->stream>>>(op->rowptr, op->colind, op->values, op->n_local, out,
                                               invert ? 1 : 0, min_diag);
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

int launch_fill(Ctx* ctx, double* y, int64_t ld, int64_t n, int k, double a) {
  if (n <= 0 || k <= 0) return 0;
  ctx->t_begin(K_AXPBY, 8.0 * n * k);
  fill_kernel<<<ew_grid(n, k), kThreads, 0, ctx->stream>>>(y, ld, (n + 3) / 4, a);
  ctx->t_end();
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

int launch_axpby(Ctx* ctx, double* y, int64_t ldy, const double* a, int64_t lda, const double* b,
                 int64_t ldb, int64_t n, int k, double alpha, double beta) {
  if (n <= 0 || k <= 0) return 0;
  const dim3 g = ew_grid(n, k);
  if (b == nullptr) {
    ctx->t_begin(K_AXPBY, 16.0 * n * k);
    axpby_kernel<0><<<g, kThreads, 0, ctx->stream>>>(y, ldy, a, lda, nullptr, 0, (n + 3) / 4,
                                                      alpha, 0.0);
  } else {
    ctx->t_begin(K_AXPBY, 24.0 * n * k);
    axpby_kernel<1><<<g, kThreads, 0, ctx->stream>>>(y, ldy, a, lda, b, ldb, (n + 3) / 4, alpha,
                                                      beta);
  }
  ctx->t_end();
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

int launch_scale_cols(Ctx* ctx, double* y, int64_t ld, int64_t n, int k,
                      const double* alpha_host) {
  if (n <= 0) return 0;
  for (int j0 = 0; j0 < k; j0 += 32) {
    const int kk = std::min(32, k - j0);
    ColScales s;
    for (int j = 0; j < kk; ++j) s.a[j] = alpha_host[j0 + j];
    ctx->t_begin(K_AXPBY, 16.0 * n * kk);
    scale_cols_kernel<<<ew_grid(n, kk), kThreads, 0, ctx->stream>>>(y + (size_t)j0 * ld, ld,
                                                                     (n + 3) / 4, s);
    ctx->t_end();
    ++ctx->launches;
    B200_CUDA(cudaGetLastError());
  }
  return 0;
}

int launch_random(Ctx* ctx, double* y, int64_t ld, int64_t n, int k, int64_t row_offset,
                  int64_t n_global, int64_t seq_start, unsigned int seed) {
  if (n <= 0 || k <= 0) return 0;
  const int64_t nchunk = (n + kRandChunk - 1) / kRandChunk;
  int64_t bx = std::min<int64_t>((nchunk + kThreads - 1) / kThreads, kNumSM * 8);
  ctx->t_begin(K_AXPBY, 8.0 * n * k);
  random_kernel<<<dim3((unsigned)std::max<int64_t>(bx, 1), (unsigned)k), kThreads, 0,
                  ctx->stream>>>(y, ld, n, row_offset, n_global, seq_start, seed);
  ctx->t_end();
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace b200
