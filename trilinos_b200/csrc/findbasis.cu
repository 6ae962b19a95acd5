// One column step of DGKS findBasis as ONE kernel.
//
// Reference call sequence per column j of a new block X (BelosDGKSOrthoManager.hpp:1000-1225):
//     oldMXj  = CloneCopy(Xj)                        :1037
//     oldDot  = MvDot(Xj, Xj)                        :1050
//     product = MvTransMv(prevX, Xj)                 :1064     prevX = X(:, 0..j-1)
//     Xj     -= prevX * product                      :1072     (MvTimesMatAddMv)
//     newDot  = MvDot(Xj, Xj)                        :1091
//     ... MvScale(Xj, 1 / sqrt(newDot))              :1217-1222
// Every operand is a single column of a block that sits in L2: each call is launch- and round-trip-bound
// (~14 us kernel + ~8 us host round trip each, 3 columns per block step at block size 4).  This kernel is the
// MvTimesMatAddMv of the sequence, and while it has the rows of the block in registers it also
//   * applies the pending normalisation factors of the previous columns (deferred MvScale, capi.cu) and writes
//     those columns back scaled -- the arithmetic the skinny Gram kernel did (a' = fl(s a), then a' enters the sums);
//   * stores the old Xj into a pending lazy copy (oldMXj);
//   * emits Xj_new . Xj_new and Xj_new . Xj_old (the MvDot calls that follow),
//   * emits Xj_new . Xk for every LATER column k of the block: with the products X0 . Xk (left behind by the block
//     update before findBasis, update.cu cross dots, for block sizes 2 and 4; by the step of column 1 otherwise),
//     these are exactly the entries of the `product` vectors of the columns still to come (scaled by the
//     normalisation factor once that is known, which the dot memo tracks), so their MvTransMv calls are answered
//     without a launch and without an exchange, and
//   * emits Xj_new . Xi for the PREVIOUS columns: the `product` of a second Gram-Schmidt pass over the same column
//     (ICGS always makes two, BelosICGSOrthoManager.hpp:1067; DGKS when its dependency test fails, :1094).
// Reference arithmetic of the update itself: Epetra_MultiVector::Multiply('N','N'), epetra/src/Epetra_MultiVector.cpp:1936-2083.
#include "device_utils.cuh"
#include "runtime.cuh"

#include <algorithm>
#include <cstdlib>

namespace b200 {

namespace {

constexpr int kFbThreads = 256;
constexpr int kFbWarps = kFbThreads / 32;
constexpr int kFbMax = 8;  // previous columns, later columns

struct FbArgs {
  double* x;
  double* A;          // previous columns (written back where a scale is pending)
  const double* next; // later columns
  double* save_old;
  int64_t lda, ldn, n;
  int p, nx, any_scale, prev0_next;
  double coef[kFbMax], scale[kFbMax];
  double alpha, beta;
  // out[0] = new.new, out[1] = new.old, out[2 + k] = new.next_k, out[2 + nx + i] = new.prev_i,
  // prev0_next: out[2 + nx + p + k] = prev_0.next_k
  FusedDots fd;
};

template <int PMAX, int NXMAX>
__global__ void __launch_bounds__(kFbThreads, PMAX <= 4 ? 2 : 1)
    fb_step_kernel(const __grid_constant__ FbArgs f) {
  constexpr int E = 2 + NXMAX + PMAX + NXMAX;
  const int64_t nvec = (f.n + 3) >> 2;
  double acc[E];  // (one accumulator per sum: the 4 rows of a lane's vector are folded pairwise before they enter it)
#pragma unroll
  for (int e = 0; e < E; ++e) acc[e] = 0.0;
  for (int64_t i = blockIdx.x * (int64_t)kFbThreads + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * kFbThreads) {
    const int64_t r = 4 * i;
    double4v va[PMAX], vn[NXMAX];
    const double4v v = ld_plain4(f.x + r);
#pragma unroll
    for (int j = 0; j < PMAX; ++j)
      if (j < f.p) va[j] = f.any_scale ? ld_plain4(f.A + (size_t)j * f.lda + r) : ld_keep4(f.A + (size_t)j * f.lda + r);
#pragma unroll
    for (int k = 0; k < NXMAX; ++k)
      if (k < f.nx) vn[k] = ld_keep4(f.next + (size_t)k * f.ldn + r);
    if (f.save_old) st_plain4(f.save_old + r, v);
    double4v s{0.0, 0.0, 0.0, 0.0};
#pragma unroll
    for (int j = 0; j < PMAX; ++j)
      if (j < f.p) {
        if (f.any_scale && f.scale[j] != 1.0) {
          va[j].x *= f.scale[j]; va[j].y *= f.scale[j]; va[j].z *= f.scale[j]; va[j].w *= f.scale[j];
          st_plain4(f.A + (size_t)j * f.lda + r, va[j]);
        }
        s.x = fma(va[j].x, f.coef[j], s.x); s.y = fma(va[j].y, f.coef[j], s.y);
        s.z = fma(va[j].z, f.coef[j], s.z); s.w = fma(va[j].w, f.coef[j], s.w);
      }
    double4v o;
    if (f.beta == 0.0) {
      o.x = f.alpha * s.x; o.y = f.alpha * s.y; o.z = f.alpha * s.z; o.w = f.alpha * s.w;
    } else {
      o.x = fma(f.alpha, s.x, f.beta * v.x); o.y = fma(f.alpha, s.y, f.beta * v.y);
      o.z = fma(f.alpha, s.z, f.beta * v.z); o.w = fma(f.alpha, s.w, f.beta * v.w);
    }
    st_plain4(f.x + r, o);
    // rows >= n of the last 4-vector are padding (arbitrary bits, possibly NaN) and must not enter the sums
    double4v vm = v;
    if (r + 3 >= f.n) {
      const bool m1 = r + 1 < f.n, m2 = r + 2 < f.n;
      if (!m1) o.y = vm.y = 0.0;
      if (!m2) o.z = vm.z = 0.0;
      o.w = vm.w = 0.0;
#pragma unroll
      for (int k = 0; k < NXMAX; ++k)
        if (k < f.nx) {
          if (!m1) vn[k].y = 0.0;
          if (!m2) vn[k].z = 0.0;
          vn[k].w = 0.0;
        }
    }
    auto dot4 = [](const double4v& a, const double4v& b) {
      return fma(a.z, b.z, a.x * b.x) + fma(a.w, b.w, a.y * b.y);
    };
    acc[0] += dot4(o, o);
    if (f.beta != 0.0) acc[1] += dot4(o, vm);
#pragma unroll
    for (int k = 0; k < NXMAX; ++k)
      if (k < f.nx) acc[2 + k] += dot4(o, vn[k]);
    // (the previous columns: zero-filled padding rows of o keep their padding out of the sums)
#pragma unroll
    for (int j = 0; j < PMAX; ++j)
      if (j < f.p) acc[2 + NXMAX + j] += dot4(o, va[j]);
    if (f.prev0_next) {
      double4v a0 = va[0];
      if (r + 3 >= f.n) {
        if (r + 1 >= f.n) a0.y = 0.0;
        if (r + 2 >= f.n) a0.z = 0.0;
        a0.w = 0.0;
      }
#pragma unroll
      for (int k = 0; k < NXMAX; ++k)
        if (k < f.nx) acc[2 + NXMAX + PMAX + k] += dot4(a0, vn[k]);
    }
  }
  __shared__ double sm[kFbWarps][E];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int e = 0; e < E; ++e) {
    const double t = warp_sum(acc[e]);
    if (lane == 0) sm[warp][e] = t;
  }
  __syncthreads();
  // compact the entries that exist: [2 + nx | p | nx (prev_0 . next)]
  const int nent = 2 + f.nx + f.p + (f.prev0_next ? f.nx : 0);
  if (threadIdx.x < nent) {
    const int e = (int)threadIdx.x;
    const int slot = e < 2 + f.nx ? e : (e < 2 + f.nx + f.p ? 2 + NXMAX + (e - 2 - f.nx) : 2 + NXMAX + PMAX + (e - 2 - f.nx - f.p));
    double t = sm[0][slot];
#pragma unroll
    for (int w = 1; w < kFbWarps; ++w) t += sm[w][slot];
    f.fd.ws[(size_t)e * gridDim.x + blockIdx.x] = t;
  }
  if (!fold_partials_wide<kFbWarps>(f.fd.ws, f.fd.ticket, nent, gridDim.x, blockIdx.x, 1, f.fd.out)) return;
  publish_result(f.fd.pub);
}

template <int PMAX, int NXMAX>
int launch_fb_t(Ctx* ctx, const FbArgs& f) {
  const int64_t nvec = (f.n + 3) / 4;
  static int per_sm = 0;  // resident CTAs per SM of this instantiation (a property of the kernel, same on every B200)
  if (per_sm == 0) {
    B200_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fb_step_kernel<PMAX, NXMAX>, kFbThreads, 0));
    per_sm = std::max(per_sm, 1);
  }
  // one wave of resident CTAs (grid-stride loop inside): no second wave whose CTAs start when the first ones fold
  const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((nvec + 2 * kFbThreads - 1) / (2 * kFbThreads),
                                                                 (int64_t)ctx->num_sms * per_sm));
  fb_step_kernel<PMAX, NXMAX><<<grid, kFbThreads, 0, ctx->stream>>>(f);
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace

bool fb_step_enabled() {
  static const bool on = [] {
    const char* e = getenv("B200_FB_STEP");
    return !(e && e[0] == '0');
  }();
  return on;
}

// x <- beta x + alpha A (scale .* coef), A = p <= 8 columns (leading dimension lda) whose pending factors
// scale[0..p) are applied and written back; next = nx <= 7 later columns.  fd.out: 2 + nx + p (+ nx) sums (see FbArgs).
int launch_fb_step(Ctx* ctx, double* x, double* A, int64_t lda, int p, const double* coef, const double* scale,
                   const double* next, int64_t ldn, int nx, bool prev0_next, int64_t n, double alpha, double beta,
                   double* save_old, const FusedDots& fd) {
  if (p < 1 || p > kFbMax || nx < 0 || nx >= kFbMax)
    return fail(B200_ERR_ARG, "MvTimesMatAddMv: findBasis step with more than 8 columns (internal)");
  FbArgs f;
  f.x = x; f.A = A; f.next = next; f.save_old = save_old; f.lda = lda; f.ldn = ldn; f.n = n; f.p = p; f.nx = nx;
  f.any_scale = 0;
  f.prev0_next = prev0_next ? 1 : 0;
  for (int j = 0; j < kFbMax; ++j) {
    f.coef[j] = j < p ? coef[j] : 0.0;
    f.scale[j] = j < p ? scale[j] : 1.0;
    if (f.scale[j] != 1.0) f.any_scale = 1;
  }
  f.alpha = alpha; f.beta = beta; f.fd = fd;
  ctx->t_begin(K_FBSTEP, 8.0 * n * ((double)p + (beta == 0.0 ? 1.0 : 2.0)));
  int rc;
  if (p <= 4 && nx <= 3) rc = launch_fb_t<4, 3>(ctx, f);
  else rc = launch_fb_t<8, 7>(ctx, f);
  ctx->t_end();
  return rc;
}

}  // namespace b200
