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

#include <cstdlib>

namespace b200 {

namespace {

constexpr int kThreadsC = 256;      // consumer threads per CTA
constexpr int kNWC = kThreadsC / 32;  // consumer warps
constexpr int kCapMax = 8184;       // largest stage capacity (nonzeros); larger tiles are chunked
constexpr int kMaxStagesSpmm = 4;
constexpr int kSlack = 96;          // stage slots beyond cap: 128-B alignment of both copies' source
                                    // (<= 31 before, <= 31 after) + batch over-read (7); keeps
                                    // (cap + kSlack) a multiple of 32 so every segment is 128-B aligned
constexpr size_t kSpmmSmemBudget = 200u << 10;  // per SM, all resident CTAs

enum SpmmMode { SPMM_PLAIN = 0, SPMM_FUSE = 1, SPMM_GHOST = 2 };

struct SpmmGeom {
  int cap;          // nonzeros per stage (multiple of 4)
  int stage_bytes;  // (cap + kSlack) * 12 + (R + 4) * 4, rounded up to 128
  int nstage;
  int chunk;        // consecutive tiles per deal (tiles are dealt to the CTAs round-robin in chunks)
};

// Persistent TMA-streamed CSR kernel.
//  * A CTA owns a contiguous range of R-row tiles (R = 256 / T), so the x gather of neighbouring
//    tiles re-uses L1/L2 lines.  One producer thread streams each tile's values / column indices /
//    row pointers -- three contiguous segments -- into a shared-memory ring with 1-D bulk copies
//    (cp.async.bulk, completion on an mbarrier), several tiles ahead of the 8 consumer warps: the
//    HBM stream never waits for the gather phase.
//  * Consumer thread t works on row r0 + (t mod R) for the K/T right-hand sides of group t / R
//    (T > 1 for long rows: a tile then still fits one stage and all 256 threads stay busy).  Each
//    row is reduced left to right, the reference's order, for every right-hand side.
//  * The gathers of U consecutive nonzeros x K/T right-hand sides (16 loads) are issued before
//    the first FMA needs one: the gather is latency-, not bandwidth-limited.
//  * Tiles with more than `cap` nonzeros take several stages ("chunks"); a thread only touches the
//    part of its row that lies in the current chunk.
// SPMM_FUSE: additionally emits dots[k] = sum y(:,k)^2 -- the MvDot(X,X) DGKS takes of the fresh
// Krylov block right after the apply (BelosDGKSOrthoManager.hpp:1269,1395).
// SPMM_GHOST: columns >= n_owned come from the halo buffer xg; with skip_ghost_rows the rows that
// touch one are left to the boundary kernel (they are computed here too, but not stored).
template <int K, int T, int MODE>
__global__ void __launch_bounds__(kThreadsC + 32, 3)
    spmm_stream_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ colind,
                       const double* __restrict__ values, int64_t n_rows, int32_t n_owned,
                       const double* __restrict__ x, int64_t ldx, const double* __restrict__ xg,
                       int64_t ldg, double* __restrict__ y, int64_t ldy, int skip_ghost_rows,
                       SpmmGeom g, const __grid_constant__ FusedDots fd) {
  constexpr int R = kThreadsC / T;   // rows per tile
  constexpr int KT = K / T;          // right-hand sides per thread
  constexpr int U = KT >= 8 ? 2 : KT >= 4 ? 4 : 8;  // nonzeros per gather batch
  constexpr int kRp = R + 4;         // row-pointer slice (R + 1 entries, 16-B multiple)
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint64_t full_bar[kMaxStagesSpmm], empty_bar[kMaxStagesSpmm];
  __shared__ double red[kNWC][KT];
  // (indexing smem_raw directly keeps the accesses in the shared address space: LDS with folded
  // immediates instead of generic loads; 16-B alignment is all the bulk copies need)
  unsigned char* const stage0 = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // Tiles are dealt to the CTAs round-robin in chunks of g.chunk consecutive tiles: the grid works inside one
  // band of rows that sweeps through the matrix, so x entries shared by tiles of neighbouring grid planes are
  // re-read out of L2 shortly after their first use instead of out of DRAM a kernel duration later (at >= 3M rows
  // x no longer fits L2), while consecutive tiles of a chunk still share L1 lines.
  // (g.chunk == 0: one contiguous, balanced range per CTA -- kept for long-row stencils, measured below.)
  const int64_t ntiles = (n_rows + R - 1) / R;
  const int64_t C = g.chunk > 0 ? g.chunk : 1;
  const int64_t nchunks = (ntiles + C - 1) / C;
  const int64_t my_chunks = (int64_t)blockIdx.x < nchunks ? (nchunks - 1 - blockIdx.x) / gridDim.x + 1 : 0;
  const int64_t t_first = ntiles * blockIdx.x / gridDim.x;
  int64_t nseq = my_chunks * C;  // tiles of this CTA
  if (my_chunks > 0 && blockIdx.x + (my_chunks - 1) * gridDim.x == nchunks - 1) nseq -= nchunks * C - ntiles;
  if (g.chunk == 0) nseq = ntiles * (blockIdx.x + 1) / gridDim.x - t_first;
  const bool contiguous = g.chunk == 0;
  auto tile_of = [&](int64_t i) {
    return contiguous ? t_first + i : ((i / C) * gridDim.x + blockIdx.x) * C + (i % C);
  };
  const uint32_t off_c = (uint32_t)(g.cap + kSlack) * 8u, off_rp = (uint32_t)(g.cap + kSlack) * 12u;

  if (threadIdx.x == 0) {
    for (int s = 0; s < g.nstage; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], kNWC);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  double ss[KT];
#pragma unroll
  for (int k = 0; k < KT; ++k) ss[k] = 0.0;

  if (warp == kNWC) {
    // ---------------------------------------------------------------- producer warp
    // all lanes fetch the segment bounds of 32 tiles at a time; lane 0 issues the copies
    int s = 0;
    uint32_t ph = 0;
    const uint64_t pol = l2_policy_evict_first();  // the matrix is streamed once per apply
    // x rows of the tile kPfDist tiles ahead (its diagonal block) are pulled into L2 while the
    // current tiles are reduced: the gather then pays L2, not HBM, latency on first touch.  All
    // CTAs advance at the same rate, so the off-diagonal neighbours (owned by other CTAs' ranges)
    // are covered by their owners' prefetches.
    constexpr int kPfDist = 4;
    const bool pf_ok = ((reinterpret_cast<uintptr_t>(x) | (uintptr_t)(ldx * 8)) & 15) == 0;
    auto prefetch_x = [&](int64_t i) {
      if (!pf_ok || i >= nseq) return;
      const int64_t r0 = tile_of(i) * R;
      const uint32_t bytes = (uint32_t)(min((int64_t)R, n_rows - r0) * 8) & ~15u;
      if (bytes == 0) return;
#pragma unroll
      for (int k = 0; k < K; ++k) bulk_prefetch_l2(x + (size_t)k * ldx + r0, bytes);
    };
    if (lane == 0)
      for (int d = 0; d < kPfDist; ++d) prefetch_x(d);
    for (int64_t tb = 0; tb < nseq; tb += 32) {
      const int64_t t = tb + lane < nseq ? tile_of(tb + lane) : 0;
      int32_t sb = 0, se = 0;
      if (tb + lane < nseq) {
        sb = __ldg(rowptr + t * R);
        se = __ldg(rowptr + min((t + 1) * (int64_t)R, n_rows));
      }
      const int nt = (int)min((int64_t)32, nseq - tb);
      for (int i = 0; i < nt; ++i) {
        const int32_t b = __shfl_sync(0xffffffffu, sb, i), e = __shfl_sync(0xffffffffu, se, i);
        if (lane == 0) {
          const int64_t r0 = tile_of(tb + i) * R;
          prefetch_x(tb + i + kPfDist);
          const uint32_t rp_cnt = (uint32_t)min((int64_t)kRp, (n_rows + 1 - r0 + 3) & ~(int64_t)3);
          int32_t base = b;
          do {
            const int32_t cnt = min(g.cap, e - base);
            // 128-B aligned sources and sizes: the bulk-copy engine moves unaligned requests in
            // 16-B pieces (measured: 8 B/clk/SM instead of the full rate)
            const int32_t g0 = base & ~31;
            const uint32_t c4 = (uint32_t)(base + cnt - g0 + 31) & ~31u;
            mbar_wait(&empty_bar[s], ph ^ 1u);
            unsigned char* st = stage0 + (size_t)s * g.stage_bytes;
            mbar_expect_tx(&full_bar[s], c4 * 12u + rp_cnt * 4u);
            if (c4) {
              bulk_load_1d(st, values + g0, c4 * 8u, &full_bar[s], pol);
              bulk_load_1d(st + off_c, colind + g0, c4 * 4u, &full_bar[s], pol);
            }
            bulk_load_1d(st + off_rp, rowptr + r0, rp_cnt * 4u, &full_bar[s], pol);
            if (++s == g.nstage) { s = 0; ph ^= 1u; }
            base += g.cap;
          } while (base < e);
        }
      }
    }
    __syncwarp();
  } else {
    // ---------------------------------------------------------------- consumer warps
    const int row_t = threadIdx.x % R;          // row of the tile
    const int kgrp = threadIdx.x / R;           // right-hand-side group (warp-uniform)
    const double* xcol[KT];                     // column bases: one IMAD.WIDE per gather
    const double* gcol[KT];
#pragma unroll
    for (int k = 0; k < KT; ++k) {
      xcol[k] = x + (size_t)(kgrp * KT + k) * ldx;
      gcol[k] = (MODE == SPMM_GHOST) ? xg + (size_t)(kgrp * KT + k) * ldg - n_owned : nullptr;
    }
    double* yk = y + (size_t)(kgrp * KT) * ldy;
    int s = 0;
    uint32_t ph = 0;
    for (int64_t seq = 0; seq < nseq; ++seq) {
      const int64_t r0 = tile_of(seq) * R;
      const int64_t r = r0 + row_t;
      const int nr = (int)min((int64_t)R, n_rows - r0);
      double acc[KT];
#pragma unroll
      for (int k = 0; k < KT; ++k) acc[k] = 0.0;
      bool has_ghost = false;
      int32_t my_b = 0, my_e = 0, base = 0, seg_e = 0;
      bool first = true;
      do {
        mbar_wait(&full_bar[s], ph);
        const unsigned char* st = stage0 + (size_t)s * g.stage_bytes;
        if (first) {
          const int32_t* srp = reinterpret_cast<const int32_t*>(st + off_rp);
          base = srp[0];
          seg_e = srp[nr];
          if (row_t < nr) {
            my_b = srp[row_t];
            my_e = srp[row_t + 1];
          }
          first = false;
        }
        const int32_t cnt = min(g.cap, seg_e - base);
        const int32_t g0 = base & ~31;
        const int32_t lo = max(my_b, base), hi = min(my_e, base + cnt);
        const double* pv = reinterpret_cast<const double*>(st) + (lo - g0);
        const int32_t* pc = reinterpret_cast<const int32_t*>(st + off_c) + (lo - g0);
        for (int32_t e = lo; e < hi; e += U, pv += U, pc += U) {
          // entries past the row end (within the stage's slack) are read but never used: their
          // column index is replaced by 0 and their FMAs are predicated off
          double v[U];
          int32_t c[U];
#pragma unroll
          for (int u = 0; u < U; ++u) {
            v[u] = pv[u];
            c[u] = (e + u < hi) ? pc[u] : 0;
          }
          double xv[U][KT];
#pragma unroll
          for (int u = 0; u < U; ++u) {
            if (MODE == SPMM_GHOST) {
              const bool gh = c[u] >= n_owned;
              has_ghost |= gh;
              // interior pass: the ghost block is being written by the neighbours right now and
              // the row is dropped anyway -- read a harmless owned entry instead
              const bool use_g = gh && !skip_ghost_rows;
              const int32_t cc = (gh && skip_ghost_rows) ? 0 : c[u];
#pragma unroll
              for (int k = 0; k < KT; ++k) xv[u][k] = ld_gather((use_g ? gcol[k] : xcol[k]) + cc);
            } else {
#pragma unroll
              for (int k = 0; k < KT; ++k) xv[u][k] = ld_gather(xcol[k] + c[u]);
            }
          }
#pragma unroll
          for (int u = 0; u < U; ++u)
#pragma unroll
            for (int k = 0; k < KT; ++k) pin(xv[u][k]);
#pragma unroll
          for (int u = 0; u < U; ++u)
            if (e + u < hi) {
#pragma unroll
              for (int k = 0; k < KT; ++k) acc[k] = fma(v[u], xv[u][k], acc[k]);
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[s]);
        if (++s == g.nstage) { s = 0; ph ^= 1u; }
        base += g.cap;
      } while (base < seg_e);
      if (row_t < nr && !(MODE == SPMM_GHOST && skip_ghost_rows && has_ghost)) {
#pragma unroll
        for (int k = 0; k < KT; ++k) {
          yk[(size_t)k * ldy + r] = acc[k];
          if (MODE == SPMM_FUSE) ss[k] = fma(acc[k], acc[k], ss[k]);
        }
      }
    }
  }
  if (MODE != SPMM_FUSE) return;
  if (warp < kNWC) {
#pragma unroll
    for (int k = 0; k < KT; ++k) {
      const double t = warp_sum(ss[k]);
      if (lane == 0) red[warp][k] = t;
    }
  }
  __syncthreads();
  if (threadIdx.x < K) {
    // right-hand side k belongs to group k / KT, i.e. to the warps [grp * R / 32, (grp + 1) * R / 32)
    const int grp = threadIdx.x / KT, kk = threadIdx.x % KT;
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < R / 32; ++w) t += red[grp * (R / 32) + w][kk];
    fd.ws[(size_t)threadIdx.x * gridDim.x + blockIdx.x] = t;
  }
  if (!fold_partials_wide<kNWC + 1>(fd.ws, fd.ticket, K, gridDim.x, blockIdx.x, 1, fd.out)) return;
  publish_result(fd.pub);
}

// boundary rows (explicit list): one thread per row, direct loads
template <int K>
__global__ void __launch_bounds__(256)
    spmm_rows_kernel(const int32_t* __restrict__ rows, int64_t nrows,
                     const int32_t* __restrict__ rowptr, const int32_t* __restrict__ colind,
                     const double* __restrict__ values, int32_t n_owned,
                     const double* __restrict__ x, int64_t ldx, const double* xg, int64_t ldg,
                     double* __restrict__ y, int64_t ldy, HaloWait w) {
  if (w.n > 0) {  // peer-memory halo: wait until every neighbour's put of this apply has landed
    if ((int)threadIdx.x < w.n) {
      spin_acquire_sys(w.flag[threadIdx.x], w.seq, w.err);  // on timeout: host sees the error word
    }
    __syncthreads();
  }
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nrows) return;
  const int32_t r = rows[i];
  double acc[K];
#pragma unroll
  for (int k = 0; k < K; ++k) acc[k] = 0.0;
  for (int32_t e = rowptr[r]; e < rowptr[r + 1]; ++e) {
    const double v = values[e];
    const int32_t c = colind[e];
    // ghost entries were written by a peer GPU (possibly during this kernel): L2-coherent loads
    if (c < n_owned) {
#pragma unroll
      for (int k = 0; k < K; ++k) acc[k] = fma(v, __ldg(x + c + (size_t)k * ldx), acc[k]);
    } else {
#pragma unroll
      for (int k = 0; k < K; ++k) acc[k] = fma(v, __ldcg(xg + (c - n_owned) + (size_t)k * ldg), acc[k]);
    }
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

// stage geometry for R-row tiles of this operator: one chunk per tile whenever two stages of the
// largest tile fit the shared-memory budget; as many resident CTAs (<= 3) and stages as fit
SpmmGeom spmm_geom(const Op* op, int R, int* ctas_per_sm) {
  const int64_t mt = op->max_tile_nnz[R == 64 ? 0 : R == 128 ? 1 : 2];
  SpmmGeom g;
  g.cap = (int)std::min<int64_t>(kCapMax, std::max<int64_t>((mt + 63) / 64 * 64, 64));
  g.stage_bytes = ((g.cap + kSlack) * 12 + (R + 4) * 4 + 127) / 128 * 128;
  static const int force_stages = [] {
    const char* e = getenv("B200_SPMM_STAGES");
    return e ? atoi(e) : 0;
  }();
  static const int force_ctas = [] {
    const char* e = getenv("B200_SPMM_CTAS");
    return e ? atoi(e) : 0;
  }();
  int ctas = 3;
  while (ctas > 1 && (size_t)ctas * (2 * g.stage_bytes + 1280) > kSpmmSmemBudget) --ctas;
  if (force_ctas > 0) ctas = std::min(ctas, force_ctas);
  int nst = (int)((kSpmmSmemBudget / ctas - 1280) / g.stage_bytes);
  nst = std::max(2, std::min(nst, kMaxStagesSpmm));
  if (force_stages >= 2) nst = std::min(nst, force_stages);
  g.nstage = nst;
  g.chunk = 1;
  *ctas_per_sm = ctas;
  return g;
}

// peer-memory halo put: dst_p[j * ld_p + i] = x[send_idx[off_p + i] + j * ldx] for every neighbour p
// (the neighbour's ghost block, over NVLink), then the last CTA raises the neighbours' flags
__global__ void __launch_bounds__(256)
    halo_put_kernel(HaloPut h, const int32_t* __restrict__ send_idx, const double* __restrict__ x,
                    int64_t ldx, uint64_t seq, unsigned int* ticket) {
  const int j = blockIdx.y;
  for (int p = 0; p < h.npeers; ++p) {
    const int32_t* idx = send_idx + h.send_off[p];
    double* dst = h.dst[p] + (size_t)j * h.dst_ld[p];
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < h.send_cnt[p];
         i += (int64_t)gridDim.x * blockDim.x)
      dst[i] = __ldg(x + (size_t)j * ldx + idx[i]);
  }
  __shared__ bool is_last;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();  // cumulative over the CTA's stores ordered by the barrier
    const unsigned int total = gridDim.x * gridDim.y;
    const unsigned int t = atomicAdd(ticket, 1u);
    is_last = (t == total - 1);
    if (is_last) *ticket = 0u;
  }
  __syncthreads();
  if (is_last && (int)threadIdx.x < h.npeers) {
    __threadfence_system();
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(h.flag[threadIdx.x]), "l"(seq) : "memory");
  }
}

template <int K, int T, int MODE>
int spmm_stream(Ctx* ctx, const Op* op, const double* x, int64_t ldx, const double* xg,
                int64_t ldg, double* y, int64_t ldy, int skip, const FusedDots& fd) {
  constexpr int R = kThreadsC / T;
  const int64_t ntiles = (op->n_local + R - 1) / R;
  if (ntiles == 0) return 0;
  static unsigned attr_devs = 0;  // bit per device: the attribute is per device, not per process
  if (!(attr_devs & (1u << (ctx->device & 31)))) {
    B200_CUDA(cudaFuncSetAttribute(spmm_stream_kernel<K, T, MODE>,
                                   cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)kSpmmSmemBudget + 128));
    attr_devs |= 1u << (ctx->device & 31);
  }
  int ctas = 1;
  SpmmGeom g = spmm_geom(op, R, &ctas);
  const int smem = g.nstage * g.stage_bytes + 128;
  const unsigned grid = (unsigned)std::min<int64_t>(ntiles, (int64_t)ctx->num_sms * ctas);
  // Measured (profiles/r2_sweep_s3.txt, cold L2): Laplace3D 200^3 k=4 300 -> 199 us and Recirc2D 2000^2 k=16
  // 499 -> 353 us with single tiles dealt round-robin (x is then read from DRAM once instead of three times);
  // Brick3D 150^3 k=8 (27 entries per row, 216 gathers per row: bound by the load pipe, not by DRAM) is fastest
  // with the contiguous ranges of round 1 (377 us vs 391).  B200_SPMM_CHUNK overrides (0: contiguous).
  static const int chunk_env = [] {
    const char* e = getenv("B200_SPMM_CHUNK");
    return e ? atoi(e) : -1;
  }();
  const double avg = op->n_local > 0 ? (double)op->nnz / (double)op->n_local : 0.0;
  g.chunk = chunk_env >= 0 ? chunk_env : (avg > 12.0 ? 0 : 1);
  spmm_stream_kernel<K, T, MODE><<<grid, kThreadsC + 32, smem, ctx->stream>>>(
      op->rowptr, op->colind, op->values, op->n_local, (int32_t)op->n_local, x, ldx, xg, ldg, y,
      ldy, skip, g, fd);
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

template <int K, int T>
int spmm_mode(Ctx* ctx, const Op* op, const double* x, int64_t ldx, const double* xg, int64_t ldg,
              double* y, int64_t ldy, int skip, const FusedDots& fd) {
  if (op->n_ghost > 0)
    return spmm_stream<K, T, SPMM_GHOST>(ctx, op, x, ldx, xg, ldg, y, ldy, skip, fd);
  if (fd.out) return spmm_stream<K, T, SPMM_FUSE>(ctx, op, x, ldx, xg, ldg, y, ldy, skip, fd);
  return spmm_stream<K, T, SPMM_PLAIN>(ctx, op, x, ldx, xg, ldg, y, ldy, skip, fd);
}

// threads per row-tile split: long rows (27-point stencils) keep all 256 consumer threads busy on
// a tile that still fits one stage by giving each of T thread groups K/T right-hand sides
int spmm_split(const Op* op, int K) {
  static const int force = [] {
    const char* e = getenv("B200_SPMM_SPLIT");
    return e ? atoi(e) : 0;
  }();
  int T = 1;
  const double avg = op->n_local > 0 ? (double)op->nnz / (double)op->n_local : 0.0;
  if (avg > 8.0 && K >= 2) T = 2;
  if (force > 0) T = force;
  while (T > 1 && (K % T != 0 || T > 4)) T >>= 1;
  if (K >= 16 && T < 2) T = 4;  // at most 8 right-hand sides per thread (register budget)
  return T;
}

template <int K>
int spmm_rows(Ctx* ctx, const Op* op, const int32_t* rows, int64_t nrows, const double* x,
              int64_t ldx, const double* xg, int64_t ldg, double* y, int64_t ldy,
              const HaloWait& w = HaloWait()) {
  if (nrows == 0) return 0;
  const unsigned grid = (unsigned)((nrows + 255) / 256);
  spmm_rows_kernel<K><<<grid, 256, 0, ctx->stream>>>(rows, nrows, op->rowptr, op->colind,
                                                      op->values, (int32_t)op->n_local, x, ldx, xg,
                                                      ldg, y, ldy, w);
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
  const int64_t ldg = op->ghost_ld;  // ghost block: one padded column per right-hand side
  int k0 = 0;
  while (k0 < k) {
    const int rem = k - k0;
    int step;
    if (rem >= 16) step = 16; else if (rem >= 8) step = 8; else if (rem >= 4) step = 4; else if (rem >= 2) step = 2; else step = 1;
    const double* xk = x + (size_t)k0 * ldx;
    const double* xgk = xg ? xg + (size_t)k0 * ldg : nullptr;
    double* yk = y + (size_t)k0 * ldy;
    int rc = 0;
    FusedDots fd;
    if (fused && fused->out && rows == nullptr && !skip && op->n_ghost == 0) {
      fd = *fused;
      fd.out = fused->out + k0;
      if (k0 + step < k) fd.not_last();
    }
    if (rows != nullptr) {
      switch (step) {
        case 16: rc = spmm_rows<16>(ctx, op, rows, nrows, xk, ldx, xgk, ldg, yk, ldy); break;
        case 8: rc = spmm_rows<8>(ctx, op, rows, nrows, xk, ldx, xgk, ldg, yk, ldy); break;
        case 4: rc = spmm_rows<4>(ctx, op, rows, nrows, xk, ldx, xgk, ldg, yk, ldy); break;
        case 2: rc = spmm_rows<2>(ctx, op, rows, nrows, xk, ldx, xgk, ldg, yk, ldy); break;
        default: rc = spmm_rows<1>(ctx, op, rows, nrows, xk, ldx, xgk, ldg, yk, ldy); break;
      }
    } else {
      const int T = spmm_split(op, step);
#define B200_SPMM_CASE(KK, TT) \
  case KK * 8 + TT: rc = spmm_mode<KK, TT>(ctx, op, xk, ldx, xgk, ldg, yk, ldy, skip, fd); break;
      switch (step * 8 + T) {
        B200_SPMM_CASE(16, 4) B200_SPMM_CASE(16, 2)
        B200_SPMM_CASE(8, 4) B200_SPMM_CASE(8, 2) B200_SPMM_CASE(8, 1)
        B200_SPMM_CASE(4, 4) B200_SPMM_CASE(4, 2) B200_SPMM_CASE(4, 1)
        B200_SPMM_CASE(2, 2) B200_SPMM_CASE(2, 1)
        B200_SPMM_CASE(1, 1)
        default: return B200_ERR_ARG;
      }
#undef B200_SPMM_CASE
    }
    if (rc) return rc;
    k0 += step;
  }
  return 0;
}

int launch_halo_put(Ctx* ctx, const HaloPut& h, const int32_t* send_idx, int64_t n_send,
                    const double* x, int64_t ldx, int k, uint64_t seq) {
  if (k <= 0) return 0;
  // few, fat CTAs: each pays one system-scope fence (remote-store acknowledgement) at its end
  const unsigned bx = (unsigned)std::max<int64_t>(1, std::min<int64_t>((n_send + 2047) / 2048, 32));
  halo_put_kernel<<<dim3(bx, (unsigned)k), 256, 0, ctx->comm_stream>>>(
      h, send_idx, x, ldx, seq, ctx->ticket + kTicketBytes / sizeof(unsigned int) - 1);
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

int launch_spmm_boundary(Ctx* ctx, const Op* op, const HaloWait& w, const double* x, int64_t ldx,
                         double* y, int64_t ldy, int k) {
  const double* xg = op->recv_buf;
  const int64_t ldg = op->ghost_ld;
  for (int k0 = 0; k0 < k;) {
    const int rem = k - k0;
    const int step = rem >= 16 ? 16 : rem >= 8 ? 8 : rem >= 4 ? 4 : rem >= 2 ? 2 : 1;
    const double* xk = x + (size_t)k0 * ldx;
    const double* xgk = xg + (size_t)k0 * ldg;
    double* yk = y + (size_t)k0 * ldy;
    int rc;
    switch (step) {
      case 16: rc = spmm_rows<16>(ctx, op, op->bnd_rows, op->n_bnd, xk, ldx, xgk, ldg, yk, ldy, w); break;
      case 8: rc = spmm_rows<8>(ctx, op, op->bnd_rows, op->n_bnd, xk, ldx, xgk, ldg, yk, ldy, w); break;
      case 4: rc = spmm_rows<4>(ctx, op, op->bnd_rows, op->n_bnd, xk, ldx, xgk, ldg, yk, ldy, w); break;
      case 2: rc = spmm_rows<2>(ctx, op, op->bnd_rows, op->n_bnd, xk, ldx, xgk, ldg, yk, ldy, w); break;
      default: rc = spmm_rows<1>(ctx, op, op->bnd_rows, op->n_bnd, xk, ldx, xgk, ldg, yk, ldy, w); break;
    }
    if (rc) return rc;
    k0 += step;
  }
  return 0;
}

int launch_pack(Ctx* ctx, cudaStream_t s, const int32_t* idx, int64_t cnt, const double* x,
                int64_t ldx, int k, double* buf) {
  if (cnt <= 0 || k <= 0) return 0;
  const unsigned bx = (unsigned)std::min<int64_t>((cnt + 255) / 256, (int64_t)ctx->num_sms * 4);
  pack_kernel<<<dim3(bx, (unsigned)k), 256, 0, s>>>(idx, cnt, x, ldx, buf);
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace b200
