// MvTransMv: C(p x q) = alpha * A^T V for a tall-skinny Krylov basis A (n x p) and a block
// V (n x q), both FP64 column-major.  This is the kernel that dominates Block GMRES
// orthogonalisation (reference: 46 % of the config-2 solve).
//
// Reference semantics: Belos::EpetraMultiVec::MvTransMv  belos/epetra/src/BelosEpetraAdapter.cpp:480-495
//   -> Epetra_MultiVector::Multiply('T','N',alpha,A,V,0)   epetra/src/Epetra_MultiVector.cpp:1936-2083
//   -> local DGEMM + Reduce() (SumAll)                     :2400-2433
//
// Three kernels share the launch interface (launch_gram): gram_tma_kernel (q <= 4 right-hand sides,
// SIMT FP64, described next), gram_dmma_kernel (q = 8 / 16, FP64 tensor cores, further down) and
// gram_skinny_kernel (one vector against <= 8 columns).
//
// Design (HBM-bound, algorithmic bytes 8 n (p + q)).  The first version of this kernel kept the
// streamed operand in registers (8 x 32-byte loads per thread per step) and reached only ~40 % of
// the HBM roofline: with ~64 KB in flight per SM the kernel is latency-, not bandwidth-bound.
// This version decouples bytes-in-flight from registers:
//   * persistent grid, one CTA per SM; CTA i owns a contiguous range of 64*SUB-row tiles and walks
//     all column blocks (TC = 8*CPW columns of A) of its rows, so every CTA streams the same number
//     of bytes and V (re-read once per column block) stays L2 resident;
//   * one elected producer thread moves each (row tile x column block) slab of A plus the matching
//     V tile into a shared-memory ring with ONE 2-D TMA tensor copy each (cp.async.bulk.tensor.2d,
//     box = R rows x TC columns; rows >= n and columns >= p are zero-filled by the TMA unit, so
//     there is no tail masking), completion counted on an mbarrier per stage; up to ~200 KB per SM
//     are in flight.  (A first TMA version issued one 1-D bulk copy per column: ptxas serialises
//     those over the lanes with an ELECT/R2UR loop, ~36 copies per 36 KB stage, and the producer
//     became the bottleneck at 53 % of the HBM roofline.)
//   * 8 consumer warps read the stage with conflict-free 128-bit LDS (lane = 2 rows), arranged as
//     WC column groups x WR row groups; a thread keeps CPW*Q FP64 accumulators for the whole row
//     range of the CTA (no per-tile reduction);
//   * per-CTA partials go to a workspace laid out [cta][entry]; the CTA that arrives last folds
//     them in a fixed order (deterministic, run-to-run reproducible) and publishes to the host
//     mailbox (or to the device buffer the NCCL allreduce works on).
#include <cuda.h>

#include "device_utils.cuh"
#include "runtime.cuh"

#include <cstdlib>

namespace b200 {

namespace {

constexpr int kMaxStages = 8;
constexpr size_t kSmemBudget = 200u << 10;
constexpr int kTC = 32;  // columns of A per column block

__device__ __forceinline__ double2 lds128(uint32_t addr) {
  double2 v;
  asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(addr));
  return v;
}
struct GramArgs {
  const double* A;
  const double* V;
  int64_t lda, ldv, n;
  int p, q;          // columns of A, columns of V handled by this launch (q <= Q)
  int tc;            // SIMT kernel: columns of A per column block (<= kTC; the blocks are made equally wide)
  int sub;           // 64-row substeps per tile (tile rows R = 64*sub)
  int rows;          // DMMA kernel: rows per tile (R = 4 mod 16: conflict-free fragment loads)
  int nstage;
  int wc_log2;       // consumer warps = WC column groups x WR row groups
  int64_t ntiles;    // ceil(n / R)
  double alpha;
  double* ws;        // [cta][p*q] partial sums
  unsigned int* ticket;
  double* out;       // C(:, 0..q) with leading dimension ldc
  int ldc;
  XchgPub pub;       // how the finishing CTA hands the result to the host
};

// Q: accumulator columns (>= q), CPW: columns of A per warp, NW: consumer warps.
// Consumer warp w = (wr, wc): column group wc (CPW columns), row group wr; a warp reads 64 rows
// (2 per lane, one conflict-free LDS.128 per column) per substep.
template <int Q, int CPW, int NW>
__global__ void __launch_bounds__((NW + 1) * 32, 1)
    gram_tma_kernel(const __grid_constant__ GramArgs g, const __grid_constant__ CUtensorMap tmA,
                    const __grid_constant__ CUtensorMap tmV) {
  constexpr int NT = (NW + 1) * 32;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint64_t full_bar[kMaxStages], empty_bar[kMaxStages];
  __shared__ double red[NW][CPW * Q];

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int R = 64 * g.sub;
  // Column blocks of equal width g.tc <= kTC (a multiple of CPW): p = 100 is 4 blocks of 28, 28, 28, 16 columns, not
  // 32, 32, 32, 4 -- a 4-column last block kept 2 of the 16 consumer warps busy for as long as a full block takes
  // (kbench p = 100: 0.86 of the HBM peak against 0.99 at p = 200).  Stage width = one block; the TMA unit zero-fills
  // columns >= p, so no lane ever reads shared memory that was not written.
  const int ncb = (g.p + g.tc - 1) / g.tc;
  const int wfull = g.tc;
  const uint32_t stage_bytes = (uint32_t)((wfull + g.q) * R) * 8u;
  unsigned char* stage0 = reinterpret_cast<unsigned char*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
  const int64_t t_begin = g.ntiles * blockIdx.x / gridDim.x;
  const int64_t t_end = g.ntiles * (blockIdx.x + 1) / gridDim.x;
  const int ntile_cta = (int)(t_end - t_begin);

  if (threadIdx.x == 0) {
    for (int s = 0; s < g.nstage; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], NW);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  if (warp == NW) {
    // ------------------------------------------------------------------ producer (one thread)
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      const uint64_t pol_a = l2_policy_evict_first();
      const uint64_t pol_v = l2_policy_evict_last();  // V is also the next kernel's operand
      for (int cb = 0; cb < ncb; ++cb) {
        const int col0 = cb * g.tc;
        for (int t = 0; t < ntile_cta; ++t) {
          mbar_wait(&empty_bar[s], ph ^ 1u);
          const int r0 = (int)((t_begin + t) * R);
          unsigned char* st = stage0 + (size_t)s * stage_bytes;
          mbar_expect_tx(&full_bar[s], stage_bytes);
          tma_load_2d(st, &tmA, r0, col0, &full_bar[s], pol_a);
          tma_load_2d(st + (size_t)wfull * R * 8, &tmV, r0, 0, &full_bar[s], pol_v);
          if (++s == g.nstage) { s = 0; ph ^= 1u; }
        }
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------------ consumer warps
    const int WC = 1 << g.wc_log2, WR = NW >> g.wc_log2;
    const int wc = warp & (WC - 1), wr = warp >> g.wc_log2;
    const uint32_t colstride = (uint32_t)R * 8u;
    const uint32_t base = smem_u32(stage0);
    const uint32_t a_off = (uint32_t)(wc * CPW) * colstride + (uint32_t)lane * 16u;
    const uint32_t v_off = (uint32_t)wfull * colstride + (uint32_t)lane * 16u;
    int s = 0;
    uint32_t ph = 0;
    for (int cb = 0; cb < ncb; ++cb) {
      const int col0 = cb * g.tc;
      const int wdt = min(g.tc, g.p - col0);
      // columns >= p of the last block are zero-filled by the TMA unit: no per-column predicate
      const bool active = wc * CPW < wdt;
      double acc[CPW][Q];
#pragma unroll
      for (int c = 0; c < CPW; ++c)
#pragma unroll
        for (int k = 0; k < Q; ++k) acc[c][k] = 0.0;

      for (int t = 0; t < ntile_cta; ++t) {
        mbar_wait(&full_bar[s], ph);
        if (active) {
          const uint32_t st = base + (uint32_t)s * stage_bytes;
          for (int ss = wr; ss < g.sub; ss += WR) {
            const uint32_t row = (uint32_t)ss * 512u;  // 64 rows x 8 bytes
            double2 a[CPW];
#pragma unroll
            for (int c = 0; c < CPW; ++c) a[c] = lds128(st + a_off + row + (uint32_t)c * colstride);
#pragma unroll
            for (int k = 0; k < Q; ++k) {
              if (k < g.q) {
                const double2 v = lds128(st + v_off + row + (uint32_t)k * colstride);
#pragma unroll
                for (int c = 0; c < CPW; ++c) {
                  acc[c][k] = fma(a[c].x, v.x, acc[c][k]);
                  acc[c][k] = fma(a[c].y, v.y, acc[c][k]);
                }
              }
            }
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[s]);
        if (++s == g.nstage) { s = 0; ph ^= 1u; }
      }

      // lanes -> warp, then the WR row-group warps of each column group, then the workspace
#pragma unroll
      for (int c = 0; c < CPW; ++c)
#pragma unroll
        for (int k = 0; k < Q; ++k) {
          const double sres = warp_sum(acc[c][k]);
          if (lane == 0) red[warp][c * Q + k] = sres;
        }
      consumer_bar<NW * 32>();
      for (int e = threadIdx.x; e < WC * CPW * Q; e += NW * 32) {
        const int grp = e / (CPW * Q), ck = e % (CPW * Q);
        const int c = ck / Q, k = ck % Q;
        const int col = col0 + grp * CPW + c;
        if (grp * CPW + c < wdt && k < g.q) {
          double sres = 0.0;
          for (int w = 0; w < WR; ++w) sres += red[(w << g.wc_log2) + grp][ck];
          g.ws[(size_t)blockIdx.x * (size_t)(g.p * g.q) + (size_t)k * g.p + col] = sres;
        }
      }
      consumer_bar<NW * 32>();
    }
  }

  // deterministic two-level fold over the per-CTA partials (thread = entry)
  const int nent = g.p * g.q;
  const bool fin = fold_partials_tall<NT>(g.ws, g.ticket, nent, (int)gridDim.x, (int)blockIdx.x,
                                          [&](int e, double sres) {
                                            const int k = e / g.p, col = e % g.p;
                                            g.out[(size_t)k * g.ldc + col] = g.alpha * sres;
                                          });
  if (!fin) return;
  publish_result(g.pub);
}

// ---------------------------------------------------------------------------------------------
// Skinny variant: one vector against p <= 8 columns.  DGKS' findBasis orthogonalises column j of
// the new block against columns 0..j-1 one column at a time (BelosDGKSOrthoManager.hpp:1345-1480):
// 3 such calls per config-2 iteration, each over a few L2-resident columns.  The TMA ring moves
// boxes one column wide for them (2 KB per copy, ~18 us per call); here every thread simply has
// all its 256-bit loads in flight at once (~5 us).
// Pending column scale factors (deferred MvScale, capi.cu): applied on the fly and written back, so the
// normalisation at the end of a findBasis column (BelosDGKSOrthoManager.hpp:1217-1222) costs no launch of its own.
struct SkinnyScale {
  double a[8];
  double v;
  int any;
};

template <int PMAX>
__global__ void __launch_bounds__(256)
    gram_skinny_kernel(const double* A, int64_t lda, int p, const double* v,
                       int64_t n, double alpha, double* __restrict__ ws, unsigned int* ticket,
                       double* __restrict__ out, const __grid_constant__ XchgPub pub,
                       const __grid_constant__ SkinnyScale sc) {
  constexpr int kW = 8;
  const int64_t nvec = (n + 3) >> 2;
  double acc[PMAX][2];
#pragma unroll
  for (int j = 0; j < PMAX; ++j) acc[j][0] = acc[j][1] = 0.0;
  for (int64_t i = blockIdx.x * 256ll + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * 256) {
    const int64_t r = 4 * i;
    double4v vb, va[PMAX];
    if (!sc.any) {
      vb = ld_keep4(v + r);
#pragma unroll
      for (int j = 0; j < PMAX; ++j)
        if (j < p) va[j] = ld_keep4(A + (size_t)j * lda + r);
    } else {
      // (coherent loads: these columns are written below; every element is read and written by this thread only)
      vb = ld_plain4(v + r);
#pragma unroll
      for (int j = 0; j < PMAX; ++j)
        if (j < p) va[j] = ld_plain4(A + (size_t)j * lda + r);
      if (sc.v != 1.0) {
        vb.x *= sc.v; vb.y *= sc.v; vb.z *= sc.v; vb.w *= sc.v;
        st_plain4(const_cast<double*>(v) + r, vb);
      }
#pragma unroll
      for (int j = 0; j < PMAX; ++j)
        if (j < p && sc.a[j] != 1.0) {
          va[j].x *= sc.a[j]; va[j].y *= sc.a[j]; va[j].z *= sc.a[j]; va[j].w *= sc.a[j];
          st_plain4(const_cast<double*>(A) + (size_t)j * lda + r, va[j]);
        }
    }
    const bool m1 = r + 1 < n, m2 = r + 2 < n, m3 = r + 3 < n;  // padding rows of the last vector
    if (!m3) {
      if (!m1) vb.y = 0.0;
      if (!m2) vb.z = 0.0;
      vb.w = 0.0;
    }
#pragma unroll
    for (int j = 0; j < PMAX; ++j)
      if (j < p) {
        if (!m3) {
          if (!m1) va[j].y = 0.0;
          if (!m2) va[j].z = 0.0;
          va[j].w = 0.0;
        }
        acc[j][0] = fma(va[j].x, vb.x, acc[j][0]);
        acc[j][1] = fma(va[j].y, vb.y, acc[j][1]);
        acc[j][0] = fma(va[j].z, vb.z, acc[j][0]);
        acc[j][1] = fma(va[j].w, vb.w, acc[j][1]);
      }
  }
  __shared__ double sm[kW][PMAX];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int j = 0; j < PMAX; ++j) {
    const double t = warp_sum(acc[j][0] + acc[j][1]);
    if (lane == 0) sm[warp][j] = t;
  }
  __syncthreads();
  if (threadIdx.x < p) {
    double t = sm[0][threadIdx.x];
#pragma unroll
    for (int w = 1; w < kW; ++w) t += sm[w][threadIdx.x];
    ws[(size_t)threadIdx.x * gridDim.x + blockIdx.x] = alpha * t;
  }
  if (!fold_partials_wide<kW>(ws, ticket, p, gridDim.x, blockIdx.x, 1, out)) return;
  publish_result(pub);
}

// ---------------------------------------------------------------------------------------------
// FP64 tensor-core (DMMA) variant for q = 8 / 16 right-hand sides.
//
// At q = 16 the Gram needs 2 q / 8 = 4 flop per streamed byte, i.e. ~26 TFLOP/s to stay on the HBM
// roofline -- 70 % of the FP64 peak (37 TFLOP/s, profiles/r1_dmma_probe_s30.txt).  The SIMT kernel
// above re-reads the V tile from shared memory for every 4 columns of A (20 LDS.128 per 128 DFMA)
// and is bound by the shared-memory pipe (34-53 % of HBM peak).  mma.sync.m8n8k4.f64 computes an
// 8 x 8 block of A^T V over 4 rows per instruction with ONE 8-byte operand per lane for A and for
// V: the same FP64 peak at 1/8 of the issue slots and ~1/10 of the shared-memory traffic.
//
// Row-tile-major: for each R-row tile the V tile is staged ONCE (its own double buffer) and the
// 32-column blocks of A stream through the stage ring against it -- a DMMA accumulator is 2
// doubles per 8 x 8 block, so a warp keeps the accumulators of all (up to 8) column blocks of a
// pass in registers (12 consumer warps + the producer = 13 warps, at most 4 per SM sub-partition,
// so each thread may use 128 registers).  Shared-memory fill traffic == algorithmic bytes, and the TMA rows stay long
// (R * 8 >= 1 KB).  12 consumer warps = 4 column tiles (8 columns of A) x 3 row groups (a narrower
// last block gets fewer column tiles and more row groups); fragment
// lane l = (row 4 step + l%4, column 8 tile + l/4); R = 4 (mod 16) makes the column stride 8 banks
// (mod 32): every fragment load is two conflict-free wavefronts.
constexpr int kMaxCB = 8;   // 32-column blocks per pass (256 columns)

template <int NTQ>  // q == 8 * NTQ
__global__ void __launch_bounds__(13 * 32, 1)
    gram_dmma_kernel(const __grid_constant__ GramArgs g, const __grid_constant__ CUtensorMap tmA,
                     const __grid_constant__ CUtensorMap tmV) {
  constexpr int NW = 12, NT = (NW + 1) * 32, Q = 8 * NTQ;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint64_t full_bar[kMaxStages], empty_bar[kMaxStages], fullv_bar[kMaxStages],
      emptyv_bar[kMaxStages];
  __shared__ double red[4 * kTC * Q];  // [row group][column of the block][rhs]

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int R = g.rows;
  const int ncb = (g.p + kTC - 1) / kTC;
  const int npass = (ncb + kMaxCB - 1) / kMaxCB;
  const int wfull = min((g.p + 7) / 8 * 8, kTC);
  const uint32_t a_bytes = (uint32_t)(wfull * R) * 8u, v_bytes = (uint32_t)(Q * R) * 8u;
  unsigned char* vbuf0 = reinterpret_cast<unsigned char*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
  const int nv = g.sub;  // V buffers: 2 when several column blocks share a tile, else ring depth
  unsigned char* stage0 = vbuf0 + (size_t)nv * v_bytes;
  const int64_t t_begin = g.ntiles * blockIdx.x / gridDim.x;
  const int64_t t_end = g.ntiles * (blockIdx.x + 1) / gridDim.x;
  const int ntile_cta = (int)(t_end - t_begin);

  if (threadIdx.x == 0) {
    for (int s = 0; s < g.nstage; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], NW);
    }
    for (int s = 0; s < nv; ++s) {
      mbar_init(&fullv_bar[s], 1);
      mbar_init(&emptyv_bar[s], NW);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  if (warp == NW) {
    if (lane == 0) {
      int s = 0, vb = 0;
      uint32_t ph = 0, vph = 0;
      const uint64_t pol_a = l2_policy_evict_first();
      const uint64_t pol_v = l2_policy_evict_last();
      for (int ps = 0; ps < npass; ++ps) {
        const int cb0 = ps * kMaxCB, cb1 = min(ncb, cb0 + kMaxCB);
        for (int t = 0; t < ntile_cta; ++t) {
          const int r0 = (int)((t_begin + t) * R);
          mbar_wait(&emptyv_bar[vb], vph ^ 1u);
          mbar_expect_tx(&fullv_bar[vb], v_bytes);
          tma_load_2d(vbuf0 + (size_t)vb * v_bytes, &tmV, r0, 0, &fullv_bar[vb], pol_v);
          if (++vb == nv) { vb = 0; vph ^= 1u; }
          for (int cb = cb0; cb < cb1; ++cb) {
            mbar_wait(&empty_bar[s], ph ^ 1u);
            mbar_expect_tx(&full_bar[s], a_bytes);
            tma_load_2d(stage0 + (size_t)s * a_bytes, &tmA, r0, cb * kTC, &full_bar[s], pol_a);
            if (++s == g.nstage) { s = 0; ph ^= 1u; }
          }
        }
      }
    }
    __syncwarp();
  } else {
    // per column block: ceil(width / 8) column tiles x NW / tiles row groups (4 x 3 for a full
    // block, 1 x 12 for a last block of <= 8 columns: every warp has work in every stage)
    const uint32_t colstride = (uint32_t)R * 8u;
    const uint32_t frag = (uint32_t)(lane >> 2) * colstride + (uint32_t)(lane & 3) * 8u;
    const uint32_t a_off = smem_u32(stage0) + frag;
    const uint32_t v_off = smem_u32(vbuf0) + frag;
    const int nsteps = R >> 2;
    int s = 0, vb = 0;
    uint32_t ph = 0, vph = 0;
    for (int ps = 0; ps < npass; ++ps) {
      const int cb0 = ps * kMaxCB;
      const int ncbp = min(ncb - cb0, kMaxCB);
      double c[kMaxCB][NTQ][2];
#pragma unroll
      for (int i = 0; i < kMaxCB; ++i)
#pragma unroll
        for (int nt = 0; nt < NTQ; ++nt) c[i][nt][0] = c[i][nt][1] = 0.0;
      for (int t = 0; t < ntile_cta; ++t) {
        mbar_wait(&fullv_bar[vb], vph);
        const uint32_t vst = v_off + (uint32_t)vb * v_bytes;
#pragma unroll
        for (int i = 0; i < kMaxCB; ++i) {
          if (i < ncbp) {
            mbar_wait(&full_bar[s], ph);
            // columns >= p of the last block are zero-filled by the TMA unit
            const int nmt = (min(kTC, g.p - (cb0 + i) * kTC) + 7) >> 3, nrg = NW / nmt;
            const int mt = warp % nmt, rg = warp / nmt;
            if (rg < nrg) {
              const uint32_t ast = a_off + (uint32_t)(8 * mt) * colstride + (uint32_t)s * a_bytes;
              for (int step = rg; step < nsteps; step += nrg) {
                const uint32_t row = (uint32_t)step * 32u;  // 4 rows x 8 bytes
                const double a = lds_f64(ast + row);
#pragma unroll
                for (int nt = 0; nt < NTQ; ++nt) {
                  const double b = lds_f64(vst + (uint32_t)(8 * nt) * colstride + row);
                  asm volatile(
                      "mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                      : "+d"(c[i][nt][0]), "+d"(c[i][nt][1])
                      : "d"(a), "d"(b));
                }
              }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty_bar[s]);
            if (++s == g.nstage) { s = 0; ph ^= 1u; }
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&emptyv_bar[vb]);
        if (++vb == nv) { vb = 0; vph ^= 1u; }
      }
      // C fragment: lane l holds C(m = l/4, n = 8 nt + 2 (l%4) + {0,1}); the 4 row groups of a
      // column tile are summed through shared memory, block by block
#pragma unroll
      for (int i = 0; i < kMaxCB; ++i) {
        if (i < ncbp) {
          const int col0 = (cb0 + i) * kTC;
          const int wdt = min(kTC, g.p - col0);
          const int nmt = (wdt + 7) >> 3, nrg = NW / nmt, wb = 8 * nmt;
          const int mt = warp % nmt, rg = warp / nmt;
          if (rg < nrg) {
            double* r = red + ((size_t)rg * wb + 8 * mt + (lane >> 2)) * Q + 2 * (lane & 3);
#pragma unroll
            for (int nt = 0; nt < NTQ; ++nt) {
              r[8 * nt] = c[i][nt][0];
              r[8 * nt + 1] = c[i][nt][1];
            }
          }
          consumer_bar<NW * 32>();
          for (int e = threadIdx.x; e < wdt * Q; e += NW * 32) {
            const int cc = e / Q, k = e % Q;
            double sres = 0.0;
            for (int r = 0; r < nrg; ++r) sres += red[((size_t)r * wb + cc) * Q + k];
            g.ws[(size_t)blockIdx.x * (size_t)(g.p * g.q) + (size_t)k * g.p + col0 + cc] = sres;
          }
          consumer_bar<NW * 32>();
        }
      }
    }
  }

  const int nent = g.p * g.q;
  const bool fin = fold_partials_tall<NT>(g.ws, g.ticket, nent, (int)gridDim.x, (int)blockIdx.x,
                                          [&](int e, double sres) {
                                            const int k = e / g.p, col = e % g.p;
                                            g.out[(size_t)k * g.ldc + col] = g.alpha * sres;
                                          });
  if (!fin) return;
  publish_result(g.pub);
}

}  // namespace

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) ==
            cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// n x ncols FP64 column-major matrix (leading dimension ld) as a 2-D tensor {row, column};
// box = rows x cols; out-of-bounds elements read as zero
int make_tmap(CUtensorMap* tm, const double* base, int64_t n, int ncols, int64_t ld, int box_rows,
              int box_cols) {
  // The encode is a pure function of its arguments and costs a driver call on the critical path of
  // every Gram / update (the host waits for the result right after): keep the last few descriptors.
  struct Entry {
    const double* base = nullptr;
    int64_t n = 0, ld = 0;
    int ncols = 0, box_rows = 0, box_cols = 0;
    CUtensorMap tm;
  };
  constexpr int kSlots = 256;
  static thread_local Entry cache[kSlots];
  uint64_t h = reinterpret_cast<uint64_t>(base) >> 8;
  h = (h ^ (uint64_t)ncols * 0x9E3779B97F4A7C15ull ^ (uint64_t)box_rows * 0xC2B2AE3D27D4EB4Full ^
       (uint64_t)box_cols * 0x165667B19E3779F9ull);
  Entry& e = cache[(h ^ (h >> 17) ^ (h >> 31)) % kSlots];
  if (e.base == base && e.n == n && e.ld == ld && e.ncols == ncols && e.box_rows == box_rows &&
      e.box_cols == box_cols) {
    *tm = e.tm;
    return 0;
  }
  EncodeTiledFn enc = encode_tiled_fn();
  if (!enc) return fail(B200_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
  const cuuint64_t dims[2] = {(cuuint64_t)n, (cuuint64_t)ncols};
  const cuuint64_t strides[1] = {(cuuint64_t)ld * sizeof(double)};
  const cuuint32_t box[2] = {(cuuint32_t)box_rows, (cuuint32_t)box_cols};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, const_cast<double*>(base), dims,
                         strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return failf(B200_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  e.base = base; e.n = n; e.ld = ld; e.ncols = ncols; e.box_rows = box_rows; e.box_cols = box_cols;
  e.tm = *tm;
  return 0;
}

namespace {

template <int Q, int CPW, int NW>
int launch_gram_t(Ctx* ctx, const double* A, int64_t lda, int p, const double* V, int64_t ldv, int q,
                  int64_t n, double alpha, double* out, int ldc, const XchgPub& pub) {
  static_assert(kTC % CPW == 0 && NW >= kTC / CPW, "one column block must fit the consumer warps");
  GramArgs g;
  g.A = A; g.V = V; g.lda = lda; g.ldv = ldv; g.n = n; g.p = p; g.q = q;
  // equally wide column blocks: ceil(p / 32) of them, width rounded up to whole warp column groups
  const int ncb = (p + kTC - 1) / kTC;
  const int wfull = std::min(((p + ncb - 1) / ncb + CPW - 1) / CPW * CPW, kTC);
  g.tc = wfull;
  // warp arrangement: as many column groups as one column block needs (power of two)
  int wc_log2 = 0;
  while ((1 << wc_log2) * CPW < wfull) ++wc_log2;
  const int WR = NW >> wc_log2;
  g.wc_log2 = wc_log2;
  g.sub = std::min(4, std::max(2, WR));  // TMA box rows <= 256
  const int R = 64 * g.sub;
  const size_t stage_bytes = (size_t)(wfull + q) * R * sizeof(double);
  g.nstage = (int)std::max<size_t>(2, std::min<size_t>(kMaxStages, kSmemBudget / stage_bytes));
  g.ntiles = (n + R - 1) / R;
  g.alpha = alpha;
  g.ws = ctx->ws; g.ticket = ctx->ticket; g.out = out; g.ldc = ldc; g.pub = pub;
  const int grid = (int)std::max<int64_t>(1, std::min<int64_t>(ctx->num_sms, g.ntiles));
  if ((size_t)p * q * (grid + grid / 16 + 1) * sizeof(double) > kWorkspaceBytes)
    return fail(B200_ERR_ARG, "MvTransMv: partial-sum workspace too small");
  CUtensorMap tmA, tmV;
  B200_TRY(make_tmap(&tmA, A, n, p, lda, R, wfull));
  B200_TRY(make_tmap(&tmV, V, n, q, ldv, R, q));
  const size_t smem = stage_bytes * g.nstage + 128;
  static unsigned attr_devs = 0;  // bit per device: the attribute is per device, not per process
  if (!(attr_devs & (1u << (ctx->device & 31)))) {
    B200_CUDA(cudaFuncSetAttribute(gram_tma_kernel<Q, CPW, NW>,
                                   cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(216u << 10)));
    attr_devs |= 1u << (ctx->device & 31);
  }
  gram_tma_kernel<Q, CPW, NW><<<grid, (NW + 1) * 32, smem, ctx->stream>>>(g, tmA, tmV);
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

template <int NTQ>
int launch_gram_dmma(Ctx* ctx, const double* A, int64_t lda, int p, const double* V, int64_t ldv,
                     int64_t n, double alpha, double* out, int ldc, const XchgPub& pub) {
  constexpr int Q = 8 * NTQ;
  GramArgs g;
  g.A = A; g.V = V; g.lda = lda; g.ldv = ldv; g.n = n; g.p = p; g.q = Q;
  const int wfull = std::min((p + 7) / 8 * 8, kTC);
  // rows per tile: R = 4 (mod 16).  One column block: V and A tiles travel together, >= 4 deep.
  // Several column blocks: two V buffers (a V tile lives for ncb stages) + >= 4 A stages.
  const bool single = p <= kTC;
  int R = (int)(kSmemBudget / ((size_t)(single ? 4 * (wfull + Q) : 4 * wfull + 2 * Q) * sizeof(double)));
  R = std::min(R, 244);
  R = (R - 4) / 16 * 16 + 4;
  g.rows = R;
  g.wc_log2 = 0;
  const size_t a_bytes = (size_t)wfull * R * sizeof(double), v_bytes = (size_t)Q * R * sizeof(double);
  if (single) {
    g.nstage = (int)std::min<size_t>(kMaxStages, kSmemBudget / (a_bytes + v_bytes));
    g.sub = g.nstage;
  } else {
    g.sub = 2;
    g.nstage = (int)std::min<size_t>(kMaxStages, (kSmemBudget - 2 * v_bytes) / a_bytes);
  }
  g.ntiles = (n + R - 1) / R;
  g.alpha = alpha;
  g.ws = ctx->ws; g.ticket = ctx->ticket; g.out = out; g.ldc = ldc; g.pub = pub;
  const int grid = (int)std::max<int64_t>(1, std::min<int64_t>(ctx->num_sms, g.ntiles));
  if ((size_t)p * Q * (grid + grid / 16 + 1) * sizeof(double) > kWorkspaceBytes)
    return fail(B200_ERR_ARG, "MvTransMv: partial-sum workspace too small");
  CUtensorMap tmA, tmV;
  B200_TRY(make_tmap(&tmA, A, n, p, lda, R, wfull));
  B200_TRY(make_tmap(&tmV, V, n, Q, ldv, R, Q));
  const size_t smem = (size_t)g.sub * v_bytes + a_bytes * g.nstage + 128;
  static unsigned attr_devs = 0;  // bit per device: the attribute is per device, not per process
  if (!(attr_devs & (1u << (ctx->device & 31)))) {
    B200_CUDA(cudaFuncSetAttribute(gram_dmma_kernel<NTQ>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)(208u << 10)));
    attr_devs |= 1u << (ctx->device & 31);
  }
  gram_dmma_kernel<NTQ><<<grid, 13 * 32, smem, ctx->stream>>>(g, tmA, tmV);
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

template <int PMAX>
int launch_gram_skinny(Ctx* ctx, const double* A, int64_t lda, int p, const double* v, int64_t n,
                       double alpha, double* out, const XchgPub& pub, const double* scales) {
  const int64_t nvec = (n + 3) / 4;
  const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((nvec + 511) / 512, ctx->num_sms * 4));
  SkinnyScale sc;
  sc.any = 0;
  sc.v = 1.0;
  for (int j = 0; j < 8; ++j) sc.a[j] = 1.0;
  if (scales) {
    for (int j = 0; j < p; ++j) sc.a[j] = scales[j];
    sc.v = scales[p];
    sc.any = 1;
  }
  gram_skinny_kernel<PMAX><<<grid, 256, 0, ctx->stream>>>(A, lda, p, v, n, alpha, ctx->ws, ctx->ticket,
                                                          out, pub, sc);
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace

static bool skinny_enabled() {
  static const bool on = [] {
    const char* e = getenv("B200_SKINNY");
    return !(e && e[0] == '0');
  }();
  return on;
}

bool gram_is_skinny(int p, int q) { return q == 1 && p >= 1 && p <= 8 && skinny_enabled(); }

int launch_gram(Ctx* ctx, const double* A, int64_t lda, int p, const double* V, int64_t ldv, int q,
                int64_t n, double alpha, int out_off, int ldc, bool last, const double* scales) {
  if (p <= 0 || q <= 0) return 0;
  if (scales && !gram_is_skinny(p, q))
    return fail(B200_ERR_ARG, "MvTransMv: pending column scales reached a kernel that cannot apply them (internal)");
  double* out = ctx->result_ptr() + out_off;
  ctx->t_begin(K_GRAM, 8.0 * n * ((double)p + q));
  // q <= 8 and q = 16 are one pass over A (Q = next supported width; unused accumulator columns
  // are masked)
  int k0 = 0;
  int rc = 0;
  while (k0 < q && rc == 0) {
    const double* Vk = V + (size_t)k0 * ldv;
    double* outk = out + (size_t)k0 * ldc;
    static const bool dmma_on = [] {
      const char* e = getenv("B200_GRAM_DMMA");
      return !(e && e[0] == '0');
    }();
    // 16 and 8 columns go to the DMMA kernel; 9..15 are split 8 + rest (two passes over A beat the
    // 16-wide SIMT kernel, which spills and reaches 37 % of the HBM peak)
    const int rem = q - k0;
    const int step = rem >= 16 ? 16 : (rem > 8 && dmma_on ? 8 : rem);
    const XchgPub flag = ctx->make_pub(last && k0 + step >= q);
    if (gram_is_skinny(p, q)) {
      if (p <= 2) rc = launch_gram_skinny<2>(ctx, A, lda, p, Vk, n, alpha, outk, flag, scales);
      else if (p <= 4) rc = launch_gram_skinny<4>(ctx, A, lda, p, Vk, n, alpha, outk, flag, scales);
      else rc = launch_gram_skinny<8>(ctx, A, lda, p, Vk, n, alpha, outk, flag, scales);
    } else if (step == 16 && dmma_on)
      rc = launch_gram_dmma<2>(ctx, A, lda, p, Vk, ldv, n, alpha, outk, ldc, flag);
    else if (step == 8 && dmma_on)
      rc = launch_gram_dmma<1>(ctx, A, lda, p, Vk, ldv, n, alpha, outk, ldc, flag);
    else if (step > 8)
      rc = launch_gram_t<16, 4, 8>(ctx, A, lda, p, Vk, ldv, step, n, alpha, outk, ldc, flag);
    else if (step > 4)
      rc = launch_gram_t<8, 4, 8>(ctx, A, lda, p, Vk, ldv, step, n, alpha, outk, ldc, flag);
    else if (step > 2)
      rc = launch_gram_t<4, 4, 16>(ctx, A, lda, p, Vk, ldv, step, n, alpha, outk, ldc, flag);
    else if (step > 1)
      rc = launch_gram_t<2, 4, 16>(ctx, A, lda, p, Vk, ldv, step, n, alpha, outk, ldc, flag);
    else
      rc = launch_gram_t<1, 4, 16>(ctx, A, lda, p, Vk, ldv, step, n, alpha, outk, ldc, flag);
    k0 += step;
  }
  ctx->t_end();
  return rc;
}

}  // namespace b200
