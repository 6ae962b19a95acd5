// Column-wise reductions with host-visible results: MvDot, MvNorm (1, 2, inf).
// Single pass over the data (256-bit loads), per-CTA partials in a workspace, the CTA that
// arrives last folds the partials in a fixed order (run-to-run reproducible) and publishes
// the k results — either straight into the host-mapped mailbox + completion flag
// (single rank) or into a device buffer that an NCCL allreduce then combines.
//
// Reference semantics: Epetra_MultiVector::Dot   epetra/src/Epetra_MultiVector.cpp:1123-1156
//                      Norm1 :1508-1545, Norm2 :1547-1577 (plain sum of squares, no scaling),
//                      NormInf :1580-1630.  sqrt for Norm2 is applied on the host after the
//                      (all)reduction, as Epetra does after SumAll.
#include "device_utils.cuh"
#include "runtime.cuh"

namespace b200 {

namespace {

// publish_result() (device_utils.cuh) restates the window layout
static_assert(kMaxPeers == 8 && kXchgMax == 8192, "window layout");

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;

enum RedOp { RED_DOT = 0, RED_DOT_SELF = 1, RED_ABS = 2, RED_MAXABS = 3 };

template <int OP>
__global__ void __launch_bounds__(kThreads)
    colreduce_kernel(const double* __restrict__ a, int64_t lda, const double* __restrict__ b,
                     int64_t ldb, int64_t n, double* __restrict__ ws, unsigned int* ticket,
                     double* __restrict__ out, const __grid_constant__ XchgPub pub) {
  const int col = blockIdx.y;
  const double* ac = a + (size_t)col * lda;
  const double* bc = (OP == RED_DOT) ? b + (size_t)col * ldb : nullptr;
  const int64_t nvec = (n + 3) >> 2;
  double acc0 = 0.0, acc1 = 0.0;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nvec;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = 4 * i;
    double4v va = ld_keep4(ac + r);
    if (r + 3 >= n) {  // mask the padding rows of the last vector
      if (r + 1 >= n) va.y = 0.0;
      if (r + 2 >= n) va.z = 0.0;
      if (r + 3 >= n) va.w = 0.0;
    }
    if (OP == RED_DOT) {
      double4v vb = ld_keep4(bc + r);
      if (r + 3 >= n) {
        if (r + 1 >= n) vb.y = 0.0;
        if (r + 2 >= n) vb.z = 0.0;
        if (r + 3 >= n) vb.w = 0.0;
      }
      acc0 = fma(va.x, vb.x, acc0); acc1 = fma(va.y, vb.y, acc1);
      acc0 = fma(va.z, vb.z, acc0); acc1 = fma(va.w, vb.w, acc1);
    } else if (OP == RED_DOT_SELF) {
      acc0 = fma(va.x, va.x, acc0); acc1 = fma(va.y, va.y, acc1);
      acc0 = fma(va.z, va.z, acc0); acc1 = fma(va.w, va.w, acc1);
    } else if (OP == RED_ABS) {
      acc0 += fabs(va.x); acc1 += fabs(va.y); acc0 += fabs(va.z); acc1 += fabs(va.w);
    } else {
      acc0 = fmax(acc0, fmax(fabs(va.x), fabs(va.y)));
      acc1 = fmax(acc1, fmax(fabs(va.z), fabs(va.w)));
    }
  }
  __shared__ double sm[kWarps];
  double v = (OP == RED_MAXABS) ? warp_max(fmax(acc0, acc1)) : warp_sum(acc0 + acc1);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) sm[warp] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = sm[0];
#pragma unroll
    for (int w = 1; w < kWarps; ++w) t = (OP == RED_MAXABS) ? fmax(t, sm[w]) : t + sm[w];
    ws[(size_t)col * gridDim.x + blockIdx.x] = t;
  }
  if (!fold_partials_wide<kWarps, OP == RED_MAXABS>(ws, ticket, gridDim.y, gridDim.x, blockIdx.x,
                                                     gridDim.y, out))
    return;
  publish_result(pub);
}

inline dim3 red_grid(int64_t n, int k) {
  const int64_t nvec = (n + 3) / 4;
  int64_t bx = (nvec + kThreads * 2 - 1) / (kThreads * 2);  // >= 2 vectors per thread
  const int64_t cap = std::max<int64_t>(1, (int64_t)kNumSM * 4 / std::max(1, k));
  if (bx > cap) bx = cap;
  if (bx < 1) bx = 1;
  return dim3((unsigned)bx, (unsigned)k, 1);
}

}  // namespace

int launch_dot(Ctx* ctx, const double* a, int64_t lda, const double* b, int64_t ldb, int64_t n,
               int k, int out_off, bool last) {
  if (k <= 0) return 0;
  const dim3 g = red_grid(n, k);
  double* out = ctx->result_ptr() + out_off;
  const XchgPub pub = ctx->make_pub(last);
  const bool self = (a == b && lda == ldb);
  ctx->t_begin(K_REDUCE, (self ? 8.0 : 16.0) * n * k);
  if (self)
    colreduce_kernel<RED_DOT_SELF><<<g, kThreads, 0, ctx->stream>>>(
        a, lda, nullptr, 0, n, ctx->ws, ctx->ticket, out, pub);
  else
    colreduce_kernel<RED_DOT><<<g, kThreads, 0, ctx->stream>>>(a, lda, b, ldb, n, ctx->ws,
                                                                ctx->ticket, out, pub);
  ctx->t_end();
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

int launch_dot_into(Ctx* ctx, const double* a, int64_t lda, int64_t n, int k, double* out) {
  if (k <= 0) return 0;
  const dim3 g = red_grid(n, k);
  const XchgPub pub;  // single-rank style with no flag: the finishing CTA only stores the sums
  colreduce_kernel<RED_DOT_SELF><<<g, kThreads, 0, ctx->stream>>>(a, lda, nullptr, 0, n, ctx->ws, ctx->ticket, out, pub);
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

int launch_norm(Ctx* ctx, const double* a, int64_t lda, int64_t n, int k, int type, int out_off,
                bool last) {
  if (k <= 0) return 0;
  const dim3 g = red_grid(n, k);
  double* out = ctx->result_ptr() + out_off;
  const XchgPub pub = ctx->make_pub(last);
  ctx->t_begin(K_REDUCE, 8.0 * n * k);
  switch (type) {
    case B200_NORM_ONE:
      colreduce_kernel<RED_ABS><<<g, kThreads, 0, ctx->stream>>>(a, lda, nullptr, 0, n, ctx->ws,
                                                                  ctx->ticket, out, pub);
      break;
    case B200_NORM_TWO:
      colreduce_kernel<RED_DOT_SELF><<<g, kThreads, 0, ctx->stream>>>(
          a, lda, nullptr, 0, n, ctx->ws, ctx->ticket, out, pub);
      break;
    default:
      colreduce_kernel<RED_MAXABS><<<g, kThreads, 0, ctx->stream>>>(
          a, lda, nullptr, 0, n, ctx->ws, ctx->ticket, out, pub);
      break;
  }
  ctx->t_end();
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

namespace {
__global__ void publish_kernel(const double* __restrict__ src, double* __restrict__ dst, int count,
                               uint64_t* flag, uint64_t seq) {
  for (int i = threadIdx.x; i < count; i += blockDim.x) dst[i] = src[i];
  publish_flag(flag, seq);
}
}  // namespace

namespace {
// One CTA per rank: the peer exchange + rank-ordered fold + mailbox publish of publish_result()
// (device_utils.cuh) as a kernel of its own -- for the results of fused reduction epilogues, which are
// exchanged asynchronously behind the kernel that produced them.
__global__ void __launch_bounds__(256) xchg_publish_kernel(const __grid_constant__ XchgPub pub) {
  publish_result(pub);
}
}  // namespace

int launch_xchg_publish(Ctx* ctx, const double* src_d, double* dst, int count, bool is_max,
                        uint64_t seq) {
  XchgPub pb;
  pb.flag = ctx->flag_d;
  pb.seq = seq;
  pb.nranks = ctx->nranks;
  pb.rank = ctx->rank;
  pb.count = count;
  pb.is_max = is_max ? 1 : 0;
  pb.xseq = ctx->xseq + 1;
  ctx->xseq += (uint64_t)std::max(1, (count + kXchgMax - 1) / kXchgMax);
  pb.src = src_d;
  pb.dst = dst;
  pb.err = ctx->flag_d + 1;
  for (int i = 0; i < ctx->nranks; ++i) pb.win[i] = ctx->xw.win[i];
  return launch_xchg_pub(ctx, pb);
}

int launch_xchg_pub(Ctx* ctx, const XchgPub& pb) {
  xchg_publish_kernel<<<1, 256, 0, ctx->stream>>>(pb);
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

namespace {
__global__ void __launch_bounds__(256)
    zero_result_kernel(double* __restrict__ out, int count, const __grid_constant__ XchgPub pub) {
  for (int i = threadIdx.x; i < count; i += blockDim.x) out[i] = 0.0;
  publish_result(pub);
}
}  // namespace

int launch_zero_result(Ctx* ctx, double* out, int count, const XchgPub& pub) {
  zero_result_kernel<<<1, 256, 0, ctx->stream>>>(out, count, pub);
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

int launch_publish(Ctx* ctx, int count) {
  publish_kernel<<<1, 256, 0, ctx->stream>>>(ctx->red_d, ctx->mbox_d, count, ctx->flag_d,
                                             ctx->seq);
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace b200
