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
#include <cuda.h>

#include "device_utils.cuh"
#include "runtime.cuh"

#include <algorithm>
#include <cstdlib>
#include <cstring>

namespace b200 {

// gram.cu: n x ncols FP64 column-major matrix as a 2-D TMA tensor {row, column}
int make_tmap(CUtensorMap* tm, const double* base, int64_t n, int ncols, int64_t ld, int box_rows,
              int box_cols);

namespace {

constexpr int kThreads = 128;
constexpr int kPChunk = 256;  // rows of B per shared-memory stage

// FUSE: the kernel additionally emits, per output column k, dots[2k] = sum new(:,k)^2 and
// dots[2k+1] = sum new(:,k)*old(:,k) (old = V before the update; only when beta != 0) — the two
// inner products DGKS asks for right after the update (BelosDGKSOrthoManager.hpp:1091,1128,1480),
// so those MvDot calls need neither a launch nor another pass over the column.
template <int Q, int UNR, bool FUSE>
__device__ __forceinline__ void update_body(double* __restrict__ V, int64_t ldv, const double* __restrict__ A,
                                            int64_t lda, int p, int64_t n, const double* __restrict__ B, int ldb,
                                            double alpha, double beta, const FusedDots& fd,
                                            double* __restrict__ save_old) {
  __shared__ double Bs[kPChunk][Q];
  // Persistent CTAs (grid = resident capacity) walk the 512-row tiles round-robin: the fused
  // reduction's fence + ticket round trip (~2 us during which the CTA streams nothing) is paid
  // once per CTA instead of once per wave of CTAs.
  const int64_t ntiles = (n + 4 * kThreads - 1) / (4 * kThreads);
  const bool single = p <= kPChunk;  // B fits one stage: load it once for all tiles
  // cross products new_i . new_k (i < k), Q <= 4 only: the `product` entries findBasis asks for next (findbasis.cu)
  constexpr int NX = (FUSE && Q >= 2 && Q <= 4) ? Q * (Q - 1) / 2 : 1;
  double dself[Q], dold[Q], dcross[NX];
#pragma unroll
  for (int k = 0; k < Q; ++k) {
    dself[k] = 0.0;
    dold[k] = 0.0;
  }
#pragma unroll
  for (int k = 0; k < NX; ++k) dcross[k] = 0.0;
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
      double4v onew[(FUSE && Q >= 2 && Q <= 4) ? Q : 1];
#pragma unroll
      for (int k = 0; k < Q; ++k) {
        double* vp = V + (size_t)k * ldv + r;
        double4v o, v{0.0, 0.0, 0.0, 0.0};
        if (beta == 0.0) {
          o.x = alpha * acc[0][k]; o.y = alpha * acc[1][k];
          o.z = alpha * acc[2][k]; o.w = alpha * acc[3][k];
        } else {
          v = ld_plain4(vp);
          // a pending lazy copy of this column (capi.cu, Q == 1 only) gets its data here
          if (Q == 1 && save_old != nullptr) st_plain4(save_old + r, v);
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
          if (Q >= 2 && Q <= 4) {
            onew[k].x = o.x; onew[k].y = m1 ? o.y : 0.0; onew[k].z = m2 ? o.z : 0.0; onew[k].w = m3 ? o.w : 0.0;
          }
        }
      }
      if (FUSE && Q >= 2 && Q <= 4 && fd.cross) {
        int e = 0;
#pragma unroll
        for (int i = 0; i < Q; ++i)
#pragma unroll
          for (int k = i + 1; k < Q; ++k, ++e) {
            double t0 = onew[i].x * onew[k].x, t1 = onew[i].y * onew[k].y;
            t0 = fma(onew[i].z, onew[k].z, t0);
            t1 = fma(onew[i].w, onew[k].w, t1);
            dcross[e] += t0 + t1;
          }
      }
    }
  }
  if (!FUSE) return;
  // CTA partials -> workspace [entry][cta]; two-level deterministic fold, last CTA publishes
  __shared__ double red[kThreads / 32][2 * Q + NX];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nent = 2 * Q + ((Q >= 2 && Q <= 4 && fd.cross) ? NX : 0);
#pragma unroll
  for (int k = 0; k < Q; ++k) {
    const double a = warp_sum(dself[k]), b = warp_sum(dold[k]);
    if (lane == 0) {
      red[warp][2 * k] = a;
      red[warp][2 * k + 1] = b;
    }
  }
  if (nent > 2 * Q) {
#pragma unroll
    for (int k = 0; k < NX; ++k) {
      const double a = warp_sum(dcross[k]);
      if (lane == 0) red[warp][2 * Q + k] = a;
    }
  }
  __syncthreads();
  if (threadIdx.x < nent) {
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) t += red[w][threadIdx.x];
    fd.ws[(size_t)threadIdx.x * gridDim.x + blockIdx.x] = t;
  }
  if (!fold_partials_wide<kThreads / 32>(fd.ws, fd.ticket, nent, gridDim.x, blockIdx.x, 1, fd.out))
    return;
  publish_result(fd.pub);
}

template <int Q, int UNR, bool FUSE>
__global__ void __launch_bounds__(kThreads, (Q <= 4 ? 4 : 3))
    update_kernel(double* __restrict__ V, int64_t ldv, const double* __restrict__ A, int64_t lda,
                  int p, int64_t n, const double* __restrict__ B, int ldb, double alpha,
                  double beta, const __grid_constant__ FusedDots fd, double* __restrict__ save_old) {
  update_body<Q, UNR, FUSE>(V, ldv, A, lda, p, n, B, ldb, alpha, beta, fd, save_old);
}

// The same kernel with the small dense block B (p x Q, compact) as a KERNEL PARAMETER: Belos hands B over on the host,
// and a staging copy + cudaMemcpyAsync in front of every update put a DMA operation (~3 us) on the critical path of a
// call whose result the host waits for.  Up to kBParamDoubles entries travel with the launch instead.
template <int Q, int UNR, bool FUSE>
__global__ void __launch_bounds__(kThreads, (Q <= 4 ? 4 : 3))
    update_kernel_bp(double* __restrict__ V, int64_t ldv, const double* __restrict__ A, int64_t lda, int p, int64_t n,
                     double alpha, double beta, const __grid_constant__ FusedDots fd, double* __restrict__ save_old,
                     const __grid_constant__ BParam bp) {
  update_body<Q, UNR, FUSE>(V, ldv, A, lda, p, n, bp.v, p, alpha, beta, fd, save_old);
}

// ---------------------------------------------------------------------------------------------
// FP64 tensor-core (DMMA) variant for q = 8 / 16 output columns.
//
// The SIMT kernel above holds 4*Q accumulators per thread, which caps it at Q = 8: a 16-column
// block ran as two passes over A (47 % of the HBM roofline, profiles/r1_kbench_b8_b16_s29.txt).
// With mma.sync.m8n8k4.f64 an 8-row x 8-column block of the product costs 2 registers per lane, so
// one pass covers all 16 columns:  C(8 rows x 8 cols) += A(8 rows x 4 basis columns) * B(4 x 8).
//   * persistent grid, one CTA per SM, contiguous row-tile range per CTA; tile = R rows
//     (R = 8 * odd: the 4 basis columns of an A fragment then land 64 bytes apart modulo 128 and a
//     fragment load is two conflict-free wavefronts);
//   * a pack kernel first re-lays B in fragment order [column block][k][pitch], pitch = 8 (mod 16),
//     zero-padded to whole blocks; the producer thread then moves each (R rows x 16 basis columns)
//     slab of A (one 2-D TMA copy) and the matching block of B (one bulk copy, L2 resident) into
//     the stage ring -- it never waits on a load itself;
//   * 12 consumer warps own the 8-row groups {w, w+12, w+24} of the tile for all basis columns:
//     accumulators stay in registers over the whole column loop; the old V values are fetched
//     into registers before the loop (each lane reads/writes 8 rows x 64 contiguous bytes per
//     column pair: full 32-byte sectors) and the update is written straight from the fragments;
//   * FUSE: the two DGKS dot products per column, as in the SIMT kernel.
constexpr int kUCB = 16;       // basis columns per stage
constexpr int kUStages = 8;
constexpr int kUMT = 3;        // 8-row groups per consumer warp
constexpr size_t kUSmemBudget = 200u << 10;

struct UpdArgs {
  double* V;
  const double* B;   // packed: [column block][kUCB][PITCH]
  int64_t ldv, n, ntiles;
  int p, rows, nstage;
  double alpha, beta;
  FusedDots fd;
};

template <int Q, int PITCH>
__global__ void pack_b_kernel(const double* __restrict__ B, int ldb, int p, double* __restrict__ out,
                              int total) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= total) return;
  const int nn = e % PITCH, gk = e / PITCH;
  out[e] = (nn < Q && gk < p) ? B[(size_t)nn * ldb + gk] : 0.0;
}

template <int NTQ, bool FUSE>
__global__ void __launch_bounds__(13 * 32, 1)
    update_dmma_kernel(const __grid_constant__ UpdArgs u, const __grid_constant__ CUtensorMap tmA) {
  constexpr int NW = 12, Q = 8 * NTQ, PITCH = (Q == 8 ? 8 : 24);
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint64_t full_bar[kUStages], empty_bar[kUStages];
  __shared__ double red[NW][2 * Q];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int R = u.rows, nmt = R >> 3;
  const int ncb = (u.p + kUCB - 1) / kUCB;
  const uint32_t a_bytes = (uint32_t)(kUCB * R) * 8u, b_bytes = (uint32_t)(kUCB * PITCH) * 8u;
  const uint32_t st_bytes = a_bytes + b_bytes;
  unsigned char* stage0 = reinterpret_cast<unsigned char*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
  const int64_t t_begin = u.ntiles * blockIdx.x / gridDim.x;
  const int64_t t_end = u.ntiles * (blockIdx.x + 1) / gridDim.x;
  const int ntile_cta = (int)(t_end - t_begin);

  if (threadIdx.x == 0) {
    for (int s = 0; s < u.nstage; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], NW);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  double dself[NTQ][2], dold[NTQ][2];
#pragma unroll
  for (int nt = 0; nt < NTQ; ++nt) dself[nt][0] = dself[nt][1] = dold[nt][0] = dold[nt][1] = 0.0;

  if (warp == NW) {
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      const uint64_t pol_a = l2_policy_evict_first();
      const uint64_t pol_b = l2_policy_evict_last();
      for (int t = 0; t < ntile_cta; ++t) {
        const int r0 = (int)((t_begin + t) * R);
        for (int cb = 0; cb < ncb; ++cb) {
          mbar_wait(&empty_bar[s], ph ^ 1u);
          unsigned char* st = stage0 + (size_t)s * st_bytes;
          mbar_expect_tx(&full_bar[s], st_bytes);
          tma_load_2d(st, &tmA, r0, cb * kUCB, &full_bar[s], pol_a);
          bulk_load_1d(st + a_bytes, u.B + (size_t)cb * (kUCB * PITCH), b_bytes, &full_bar[s], pol_b);
          if (++s == u.nstage) { s = 0; ph ^= 1u; }
        }
      }
    }
    __syncwarp();
  } else {
    const uint32_t colstride = (uint32_t)R * 8u;
    // A fragment: lane l = (row l/4 of the 8-row group, basis column l%4 of the step)
    const uint32_t a_off = smem_u32(stage0) + (uint32_t)(lane & 3) * colstride + (uint32_t)(lane >> 2) * 8u;
    // B fragment: lane l = (k = l%4, n = l/4)
    const uint32_t b_off = smem_u32(stage0) + a_bytes + (uint32_t)((lane & 3) * PITCH + (lane >> 2)) * 8u;
    int s = 0;
    uint32_t ph = 0;
    for (int t = 0; t < ntile_cta; ++t) {
      const int64_t r0 = (t_begin + t) * R;
      double acc[kUMT][NTQ][2], vold[kUMT][NTQ][2];
      bool on[kUMT];
#pragma unroll
      for (int i = 0; i < kUMT; ++i) {
        const int m = warp + NW * i;
        const int64_t row = r0 + 8 * m + (lane >> 2);
        on[i] = m < nmt && row < u.n;
#pragma unroll
        for (int nt = 0; nt < NTQ; ++nt)
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            acc[i][nt][j] = 0.0;
            vold[i][nt][j] = 0.0;
            if (u.beta != 0.0 && on[i])
              vold[i][nt][j] = u.V[(size_t)(8 * nt + 2 * (lane & 3) + j) * u.ldv + row];
          }
      }
      for (int cb = 0; cb < ncb; ++cb) {
        mbar_wait(&full_bar[s], ph);
        const uint32_t ast = a_off + (uint32_t)s * st_bytes, bst = b_off + (uint32_t)s * st_bytes;
#pragma unroll
        for (int ks = 0; ks < kUCB / 4; ++ks) {
          double b[NTQ];
#pragma unroll
          for (int nt = 0; nt < NTQ; ++nt) b[nt] = lds_f64(bst + (uint32_t)(4 * ks * PITCH + 8 * nt) * 8u);
#pragma unroll
          for (int i = 0; i < kUMT; ++i) {
            if (warp + NW * i < nmt) {
              const double a = lds_f64(ast + (uint32_t)(4 * ks) * colstride + (uint32_t)(warp + NW * i) * 64u);
#pragma unroll
              for (int nt = 0; nt < NTQ; ++nt)
                asm volatile(
                    "mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                    : "+d"(acc[i][nt][0]), "+d"(acc[i][nt][1])
                    : "d"(a), "d"(b[nt]));
            }
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[s]);
        if (++s == u.nstage) { s = 0; ph ^= 1u; }
      }
      // C fragment: lane l holds rows l/4, columns 8 nt + 2 (l%4) + {0,1}
#pragma unroll
      for (int i = 0; i < kUMT; ++i) {
        if (on[i]) {
          const int64_t row = r0 + 8 * (warp + NW * i) + (lane >> 2);
#pragma unroll
          for (int nt = 0; nt < NTQ; ++nt)
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              const double v = vold[i][nt][j];
              const double o = u.beta == 0.0 ? u.alpha * acc[i][nt][j]
                                             : fma(u.alpha, acc[i][nt][j], u.beta * v);
              u.V[(size_t)(8 * nt + 2 * (lane & 3) + j) * u.ldv + row] = o;
              if (FUSE) {
                dself[nt][j] = fma(o, o, dself[nt][j]);
                dold[nt][j] = fma(o, v, dold[nt][j]);
              }
            }
        }
      }
    }
  }
  if (!FUSE) return;
  // lanes with equal l%4 hold the same columns: fold over l/4, then over the consumer warps
  if (warp < NW) {
#pragma unroll
    for (int nt = 0; nt < NTQ; ++nt)
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        double a = dself[nt][j], b = dold[nt][j];
#pragma unroll
        for (int o = 4; o < 32; o <<= 1) {
          a += __shfl_xor_sync(0xffffffffu, a, o);
          b += __shfl_xor_sync(0xffffffffu, b, o);
        }
        if (lane < 4) {
          const int k = 8 * nt + 2 * lane + j;
          red[warp][2 * k] = a;
          red[warp][2 * k + 1] = b;
        }
      }
  }
  __syncthreads();
  if (threadIdx.x < 2 * Q) {
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < NW; ++w) t += red[w][threadIdx.x];
    u.fd.ws[(size_t)threadIdx.x * gridDim.x + blockIdx.x] = t;
  }
  if (!fold_partials_wide<NW + 1>(u.fd.ws, u.fd.ticket, 2 * Q, gridDim.x, blockIdx.x, 1, u.fd.out))
    return;
  publish_result(u.fd.pub);
}

template <int NTQ>
int launch_update_dmma(Ctx* ctx, double* V, int64_t ldv, const double* A, int64_t lda, int p,
                       int64_t n, const double* dB, int ldb, double alpha, double beta,
                       const FusedDots& fd) {
  constexpr int Q = 8 * NTQ, PITCH = (Q == 8 ? 8 : 24);
  UpdArgs u;
  u.V = V; u.B = ctx->dBpack; u.ldv = ldv; u.n = n; u.p = p; u.alpha = alpha; u.beta = beta;
  u.fd = fd;
  const int total = (p + kUCB - 1) / kUCB * kUCB * PITCH;
  if ((size_t)total * sizeof(double) > kPackBytes)
    return fail(B200_ERR_ARG, "MvTimesMatAddMv: dense block larger than the fragment buffer");
  pack_b_kernel<Q, PITCH><<<(total + 255) / 256, 256, 0, ctx->stream>>>(dB, ldb, p, ctx->dBpack, total);
  ++ctx->launches;
  // tile rows: 8 * {11, 23, 31} (1, 2 or 2-3 row groups per consumer warp); take the one that
  // leaves the least idle tail when the tiles are dealt to the CTAs
  // (larger tiles amortise the per-stage hand-over better: a 4 % longer tail is accepted for them)
  int best = 0;
  int64_t min_cost = 0;
  const int cand[3] = {88, 184, 248};  // a TMA box is at most 256 rows
  int64_t cost[3];
  for (int i = 0; i < 3; ++i) {
    const int64_t nt = (n + cand[i] - 1) / cand[i];
    const int64_t grid = std::min<int64_t>(ctx->num_sms, nt);
    cost[i] = (nt + grid - 1) / grid * cand[i];
    if (i == 0 || cost[i] < min_cost) min_cost = cost[i];
  }
  for (int i = 0; i < 3; ++i)
    if (cost[i] * 100 <= min_cost * 104) best = cand[i];
  u.rows = best;
  u.ntiles = (n + best - 1) / best;
  const size_t st_bytes = (size_t)kUCB * best * 8 + (size_t)kUCB * PITCH * 8;
  u.nstage = (int)std::min<size_t>(kUStages, kUSmemBudget / st_bytes);
  const int grid = (int)std::max<int64_t>(1, std::min<int64_t>(ctx->num_sms, u.ntiles));
  CUtensorMap tmA;
  B200_TRY(make_tmap(&tmA, A, n, p, lda, best, kUCB));
  const size_t smem = st_bytes * u.nstage + 128;
  static unsigned attr_devs = 0;  // bit per device: the attribute is per device, not per process
  if (!(attr_devs & (1u << (ctx->device & 31)))) {
    B200_CUDA(cudaFuncSetAttribute(update_dmma_kernel<NTQ, true>,
                                   cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(208u << 10)));
    B200_CUDA(cudaFuncSetAttribute(update_dmma_kernel<NTQ, false>,
                                   cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(208u << 10)));
    attr_devs |= 1u << (ctx->device & 31);
  }
  if (fd.out)
    update_dmma_kernel<NTQ, true><<<grid, 13 * 32, smem, ctx->stream>>>(u, tmA);
  else
    update_dmma_kernel<NTQ, false><<<grid, 13 * 32, smem, ctx->stream>>>(u, tmA);
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

template <int Q, int UNR>
int launch_update_t(Ctx* ctx, double* V, int64_t ldv, const double* A, int64_t lda, int p,
                    int64_t n, const double* dB, int ldb, double alpha, double beta,
                    const FusedDots& fd, double* save_old = nullptr, const double* hB = nullptr) {
  const int64_t ntiles = (n + 4 * kThreads - 1) / (4 * kThreads);
  static int per_sm = 0;  // resident CTAs per SM of this instantiation (a property of the kernel, same on every B200)
  if (per_sm == 0) {
    B200_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, update_kernel<Q, UNR, true>,
                                                            kThreads, 0));
    per_sm = std::max(per_sm, 1);
  }
  const unsigned grid = (unsigned)std::min<int64_t>(ntiles, (int64_t)ctx->num_sms * per_sm);
  if (hB) {  // B (host, compact p x Q) travels as a kernel parameter
    static thread_local BParam bp;  // (8 KB: not on the stack of every call)
    std::memcpy(bp.v, hB, sizeof(double) * (size_t)p * Q);
    if (fd.out)
      update_kernel_bp<Q, UNR, true><<<grid, kThreads, 0, ctx->stream>>>(V, ldv, A, lda, p, n, alpha, beta, fd, save_old, bp);
    else
      update_kernel_bp<Q, UNR, false><<<grid, kThreads, 0, ctx->stream>>>(V, ldv, A, lda, p, n, alpha, beta, fd, save_old, bp);
    ++ctx->launches;
    B200_CUDA(cudaGetLastError());
    return 0;
  }
  if (fd.out)
    update_kernel<Q, UNR, true><<<grid, kThreads, 0, ctx->stream>>>(V, ldv, A, lda, p, n, dB, ldb,
                                                                     alpha, beta, fd, save_old);
  else
    update_kernel<Q, UNR, false><<<grid, kThreads, 0, ctx->stream>>>(V, ldv, A, lda, p, n, dB,
                                                                      ldb, alpha, beta, fd, save_old);
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace

int launch_update(Ctx* ctx, double* V, int64_t ldv, int q, const double* A, int64_t lda, int p,
                  int64_t n, const double* dB, int ldb, double alpha, double beta,
                  const FusedDots* fused, double* save_old, const double* hB) {
  if (n <= 0 || q <= 0) return 0;
  if (save_old && (q != 1 || beta == 0.0))
    return fail(B200_ERR_ARG, "MvTimesMatAddMv: save_old needs a one-column read-modify-write update (internal)");
  ctx->t_begin(K_UPDATE, 8.0 * n * ((double)p + (beta == 0.0 ? 1.0 : 2.0) * q));
  int k0 = 0, rc = 0;
  while (k0 < q && rc == 0) {
    const int rem = q - k0;
    double* Vk = V + (size_t)k0 * ldv;
    const double* Bk = dB + (size_t)k0 * ldb;
    static const int dmma_min = [] {  // B200_UPDATE_DMMA=0: SIMT kernels only, =16: DMMA for 16 only
      const char* e = getenv("B200_UPDATE_DMMA");
      return e ? (atoi(e) == 0 ? 1 << 30 : atoi(e)) : 8;
    }();
    int step;
    if (rem >= 16 && dmma_min <= 16) step = 16;
    else if (rem >= 8) step = 8; else if (rem >= 4) step = 4; else if (rem >= 2) step = 2; else step = 1;
    FusedDots fd;
    if (fused && fused->out) {
      fd = *fused;
      fd.out = fused->out + 2 * k0;
      if (k0 + step < q) fd.not_last();  // only the last launch hands the results over
      if (step != q || q > 4) fd.cross = 0;  // (the caller asks for cross products only where one SIMT launch covers q)
    }
    switch (step) {
      case 16: rc = launch_update_dmma<2>(ctx, Vk, ldv, A, lda, p, n, Bk, ldb, alpha, beta, fd); break;
      case 8:
        if (dmma_min <= 8) {
          rc = launch_update_dmma<1>(ctx, Vk, ldv, A, lda, p, n, Bk, ldb, alpha, beta, fd);
          break;
        }
        rc = launch_update_t<8, 4>(ctx, Vk, ldv, A, lda, p, n, Bk, ldb, alpha, beta, fd); break;
      // (hB: the compact host copy of B, when the caller did not upload it: columns k0 .. of it for this launch)
      case 4: rc = launch_update_t<4, 8>(ctx, Vk, ldv, A, lda, p, n, Bk, ldb, alpha, beta, fd, nullptr, hB ? hB + (size_t)k0 * p : nullptr); break;
      case 2: rc = launch_update_t<2, 8>(ctx, Vk, ldv, A, lda, p, n, Bk, ldb, alpha, beta, fd, nullptr, hB ? hB + (size_t)k0 * p : nullptr); break;
      default: rc = launch_update_t<1, 8>(ctx, Vk, ldv, A, lda, p, n, Bk, ldb, alpha, beta, fd, save_old, hB ? hB + (size_t)k0 * p : nullptr); break;
    }
    k0 += step;
  }
  ctx->t_end();
  return rc;
}

}  // namespace b200
