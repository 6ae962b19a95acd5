// MvTimesMatAddMv fused with the Gram product the orthogonalisation asks for next.
//
// Block Gram-Schmidt in Belos always makes two passes over the basis per block step
// (BelosDGKSOrthoManager.hpp:1400-1472, max_blk_ortho_ = 2; BelosICGSOrthoManager.hpp the same loop):
//     C  = Q^T X          (MvTransMv)           X <- X - Q C    (MvTimesMatAddMv)
//     C2 = Q^T X          (MvTransMv)           X <- X - Q C2   (MvTimesMatAddMv)
// i.e. the basis Q (n x p, the only large operand) crosses HBM four times.  The second Gram reads
// exactly the tile of Q the first update has just had in shared memory.  This kernel does the update
// and, from the same staged tile, accumulates Q^T X_new: the result waits in the host mailbox, and the
// MvTransMv call that follows is answered without a launch (capi.cu, gram memo, keyed by the content
// ids of the columns, as the fused dot products already are).  One pass over Q of four disappears.
//
// Reference semantics are those of the two separate calls:
//   MvTimesMatAddMv  belos/epetra/src/BelosEpetraAdapter.cpp:421-435 -> Epetra_MultiVector::Multiply('N','N')
//   MvTransMv        belos/epetra/src/BelosEpetraAdapter.cpp:480-495 -> Epetra_MultiVector::Multiply('T','N') + SumAll
//
// Design (HBM-bound: algorithmic bytes are those of the update alone, 8 n (p + 2 q)):
//   * persistent grid, one CTA per SM, contiguous range of R-row tiles per CTA.  A tile (R rows x all
//     p <= 256 basis columns) must be resident when X_new of its rows is known, so the ring holds 2..4
//     whole tiles (ug_bufs: what fits in 224 KB); each tile arrives as ceil(p/16) TMA boxes (R rows x 16
//     columns) with an mbarrier each, so the update starts on the first box while the rest is in flight.
//     A buffer is handed back to the producer 64 columns (kGC boxes, a "release group") at a time, as the
//     Gram sweep leaves them: the refill of the early columns overlaps the sweep of the late ones.
//   * the Gram product runs on the FP64 tensor cores (mma.sync.m8n8k4.f64): a DMMA accumulator block
//     costs 2 registers per lane, so ONE warp can hold Q^T X_new for all 256 columns (64 registers) --
//     with SIMT lanes-over-rows accumulators that would be 32x replicated and the kernel could not be
//     row-tile-major.  q <= 8 right-hand sides; unused accumulator columns are fed zeros.
//   * warp roles: 4 update warps + 8 Gram warps + 1 producer = 13 warps (at most 4 per SM sub-partition:
//     128 registers per thread).  The old X tile comes by TMA with the LAST box of its tile, in the
//     [column][R] layout X_new is left in; the update warps overwrite it in place, write X_new to global
//     memory and keep the two DGKS dot products per column (x_new.x_new, x_new.x_old) as the plain
//     update kernels do.  Update role: DMMA (update_role<K>, warp w owns the 8-row groups {w, w+4, ...})
//     for q > 4 or p > 80; SIMT DFMA (update_role_simt) otherwise, where half of every DMMA would be
//     padding (measured cross-over, launch_update_gram).  Gram warp wg takes the 4-row steps
//     g = wg (mod 8) of every tile, release group by release group (8 independent accumulator chains).
//   * R = 28 (mod 32): the column stride is 12 (mod 16) doubles, every fragment load of either phase is
//     two conflict-free wavefronts, and the 8-row groups deal evenly to the 4 update warps.
//   * per-CTA partials (p q Gram entries + 2 q dots) -> workspace; the deterministic two-level fold and
//     the hand-over to the host (or the peer exchange, multi-rank) are those of the other kernels.
//   Measured: bounded by tile residency, 0.63-0.72 of the HBM peak for its own bytes.
#include <cuda.h>

#include "device_utils.cuh"
#include "runtime.cuh"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace b200 {

int make_tmap(CUtensorMap* tm, const double* base, int64_t n, int ncols, int64_t ld, int box_rows,
              int box_cols);

namespace {

constexpr int kCW = 16;          // basis columns per TMA box
constexpr int kMaxChunks = 16;   // p <= 256
constexpr int kNU = 4, kNG = 8;  // update / Gram warps
constexpr int kMaxCB = 32;       // 8-column blocks (p <= 256)
constexpr int kMaxNB = 4;        // tile buffers
constexpr int kGC = 4;           // boxes per release group: a buffer is refilled 64 columns at a time, as the Gram sweep leaves them
constexpr int kThreadsUG = (kNU + kNG + 1) * 32;  // 13 warps, at most 4 per sub-partition: 128 registers per thread
constexpr size_t kSmemUG = 224u << 10;

struct UGArgs {
  double* V;
  const double* B;  // p x q, leading dimension ldb (device)
  int64_t ldv, n, ntiles;
  int p, q, ldb, rows, nchunk, nbuf, dmma_update;
  double alpha, beta;
  FusedDots fd;     // out[0 .. 2q): dot products, out[2q + k p + col]: (A^T V_new)(col, k)
  BParam bval;      // B == nullptr: the dense block travels with the launch (compact p x q)
};

// shared-memory geometry, the same on host and device
struct UGSmem {
  uint32_t bs_bytes, x_bytes, chunk_bytes, tile_bytes;
  B200_HD UGSmem(int R, int nchunk, int q) {
    bs_bytes = (uint32_t)nchunk * kCW * 64u;                  // B as [column][8], zero padded
    x_bytes = ((uint32_t)(q * R) * 8u + 127u) & ~127u;        // one X_new buffer
    chunk_bytes = (uint32_t)(R * kCW) * 8u;
    tile_bytes = chunk_bytes * (uint32_t)nchunk;
  }
  B200_HD size_t total(int nbuf) const { return (size_t)bs_bytes + (size_t)nbuf * (x_bytes + tile_bytes) + 128 + 64; }
};

struct UGBars {
  uint64_t full[kMaxNB][kMaxChunks], gempty[kMaxNB][kMaxChunks / kGC], xfull[kMaxNB];
};

// -DB200_UG_TRACE: CTA 0 records (globaltimer, event, tile) per warp -- the timeline behind DESIGN.md 3a
// (tools/ug_trace.py).  Not compiled into the product.
#ifdef B200_UG_TRACE
__device__ unsigned long long g_ug_trace[13][4096][2];
__device__ unsigned int g_ug_trace_n[13];
__device__ __forceinline__ void ug_trace(int warp, int lane, int ev, int tile) {
  if (blockIdx.x != 0 || lane != 0) return;
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  const unsigned int i = g_ug_trace_n[warp]++;
  if (i < 4096) {
    g_ug_trace[warp][i][0] = t;
    g_ug_trace[warp][i][1] = ((unsigned long long)ev << 32) | (unsigned int)tile;
  }
}
#define UG_TRACE(warp, lane, ev, tile) ug_trace(warp, lane, ev, tile)
#else
#define UG_TRACE(warp, lane, ev, tile)
#endif

#define B200_DMMA(c0, c1, a, b)                                                                \
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" \
               : "+d"(c0), "+d"(c1)                                                            \
               : "d"(a), "d"(b))

// ------------------------------------------------------------------ update warps
// K = 8-row groups per warp (R = 32 K - 4): every loop over the groups is unrolled without predicates.  (A first
// version unrolled to the maximum and predicated each slot on a run-time count: the chain ISETP -> @P LDS -> @P DMMA
// through ONE predicate register serialised all 8 slots, ~280 cycles per 4-column step whatever the count --
// 5x slower than the two separate kernels, profiles/r2_ug_predication_s8.txt.)
//
// One warp issues a DMMA every ~27 cycles at best (ncu: a fixed wait follows each one; the pipe of its
// sub-partition takes one every 16), and a DMMA that accumulates onto the previous one of its chain waits for
// it.  So: fragment loads in batches of K x N (one scoreboard wait per batch, not per DMMA), NCH
// interleaved accumulator chains per row group, and a single-group warp (K = 1) takes two boxes per batch.
template <int K, int N, int NCH>
__device__ __forceinline__ void update_batch(double (&acc)[K][NCH][2], uint32_t a_addr, uint32_t b_addr,
                                             uint32_t colstride, int chain0) {
  double b[N], a[N][K];
#pragma unroll
  for (int ks = 0; ks < N; ++ks) b[ks] = lds_f64(b_addr + (uint32_t)(4 * ks) * 64u);
#pragma unroll
  for (int ks = 0; ks < N; ++ks)
#pragma unroll
    for (int i = 0; i < K; ++i) a[ks][i] = lds_f64(a_addr + (uint32_t)(4 * ks) * colstride + (uint32_t)(kNU * i) * 64u);
#pragma unroll
  for (int ks = 0; ks < N; ++ks)
#pragma unroll
    for (int i = 0; i < K; ++i)
      B200_DMMA(acc[i][(chain0 + ks) % NCH][0], acc[i][(chain0 + ks) % NCH][1], a[ks][i], b[ks]);
}

template <int K>
__device__ __forceinline__ void update_role(const UGArgs& u, const UGSmem& g, UGBars& bar, unsigned char* xbuf0,
                                            unsigned char* tile0, const double* Bs, int64_t t_begin, int ntile_cta,
                                            int warp, int lane, double (*dred)[16]) {
  // boxes per loop iteration / 4-column steps per batch: 16 DMMAs per batch where the registers allow it
  constexpr int CS = K == 1 ? 4 : (K == 2 ? 2 : 1);
  constexpr int NB = K <= 4 ? 4 : 2;  // steps per batch when a single box is processed (K x NB <= 16)
  constexpr int NCH = K == 1 ? 4 : 2;
  const int R = u.rows, q = u.q, nchunk = u.nchunk;
  const uint32_t colstride = (uint32_t)R * 8u;
  // A fragment: lane l = (row l/4 of the group, basis column l%4 of the step); B: (k = l%4, n = l/4)
  const uint32_t a_lane = (uint32_t)(lane & 3) * colstride + (uint32_t)(lane >> 2) * 8u + (uint32_t)warp * 64u;
  const uint32_t b_addr = smem_u32(Bs) + (uint32_t)((lane & 3) * 8 + (lane >> 2)) * 8u;
  const int c0 = 2 * (lane & 3);  // this lane's output columns c0, c0 + 1
  double dself[2] = {0.0, 0.0}, dold[2] = {0.0, 0.0};  // dot products of this lane's two columns
  int tb = 0;
  uint32_t par = 0;
  for (int t = 0; t < ntile_cta; ++t) {
    const int64_t r0 = (t_begin + t) * R;
    double acc[K][NCH][2];
#pragma unroll
    for (int i = 0; i < K; ++i)
#pragma unroll
      for (int h = 0; h < NCH; ++h) acc[i][h][0] = acc[i][h][1] = 0.0;
    const uint32_t tile_addr = smem_u32(tile0) + (uint32_t)tb * g.tile_bytes + a_lane;
    // columns >= p of the last box are zero-filled by the TMA unit and meet zero rows of Bs: no column predicate
    // (a release group of kGC boxes is handed back as soon as this warp has read it -- except the last one, which
    // goes with the tile's X buffer at the end)
    int c = 0;
    for (; c + CS <= nchunk; c += CS) {
#pragma unroll
      for (int i = 0; i < CS; ++i) mbar_wait(&bar.full[tb][c + i], par);
      UG_TRACE(warp, lane, 10 + c, t);
#ifndef B200_UG_NO_UPDATE
      if (CS > 1) {
        update_batch<K, 4 * CS, NCH>(acc, tile_addr + (uint32_t)(kCW * c) * colstride, b_addr + (uint32_t)(kCW * c) * 64u,
                                     colstride, 0);
      } else {
#pragma unroll
        for (int h = 0; h < 4; h += NB)
          update_batch<K, NB, NCH>(acc, tile_addr + (uint32_t)(kCW * c + 4 * h) * colstride,
                                   b_addr + (uint32_t)(kCW * c + 4 * h) * 64u, colstride, 0);
      }
#endif
      if (((c + CS) & (kGC - 1)) == 0 && c + CS < nchunk) {
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar.gempty[tb][(c + CS) / kGC - 1]);
      }
    }
    for (; c < nchunk; ++c) {  // (CS > 1 only: boxes left over)
      mbar_wait(&bar.full[tb][c], par);
#ifndef B200_UG_NO_UPDATE
      update_batch<K, 4, NCH>(acc, tile_addr + (uint32_t)(kCW * c) * colstride, b_addr + (uint32_t)(kCW * c) * 64u,
                              colstride, 0);
#endif
      if (((c + 1) & (kGC - 1)) == 0 && c + 1 < nchunk) {
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar.gempty[tb][(c + 1) / kGC - 1]);
      }
    }
    // C fragment: lane l holds rows l/4, columns 2 (l%4) + {0,1}.  The old X tile came with the last box
    // (rows >= n zero-filled); X_new replaces it in place, element by element, for the Gram warps.
    double* xb = reinterpret_cast<double*>(xbuf0 + (size_t)tb * g.x_bytes);
#pragma unroll
    for (int i = 0; i < K; ++i) {
      const int rt = 8 * (warp + kNU * i) + (lane >> 2);
      const int64_t row = r0 + rt;
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        if (c0 + j < q && rt < R) {
          double s = acc[i][0][j] + acc[i][1][j];
          if (NCH == 4) s += acc[i][2][j] + acc[i][3][j];
          double o;
          if (u.beta == 0.0) {
            o = row < u.n ? u.alpha * s : 0.0;
            dself[j] = fma(o, o, dself[j]);
          } else {
            const double v = xb[(c0 + j) * R + rt];
            o = fma(u.alpha, s, u.beta * v);
            dself[j] = fma(o, o, dself[j]);
            dold[j] = fma(o, v, dold[j]);
          }
          if (row < u.n) u.V[(size_t)(c0 + j) * u.ldv + row] = o;
          xb[(c0 + j) * R + rt] = o;
        }
      }
    }
    __syncwarp();
    UG_TRACE(warp, lane, 2, t);
    if (lane == 0) {
      mbar_arrive(&bar.xfull[tb]);
      mbar_arrive(&bar.gempty[tb][(nchunk - 1) / kGC]);
    }
    if (++tb == u.nbuf) { tb = 0; par ^= 1u; }
  }
  // lanes with equal l%4 hold the same columns: fold over l/4
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    double a = dself[j], b = dold[j];
#pragma unroll
    for (int o = 4; o < 32; o <<= 1) {
      a += __shfl_xor_sync(0xffffffffu, a, o);
      b += __shfl_xor_sync(0xffffffffu, b, o);
    }
    if (lane < 4) {
      dred[warp][2 * (2 * lane + j)] = a;
      dred[warp][2 * (2 * lane + j) + 1] = b;
    }
  }
  // the arrivals that match the Gram warps' reduction rounds (gram_role)
#pragma unroll
  for (int round = 0; round <= kNG / 4; ++round) consumer_bar<(kNU + kNG) * 32>();
}

// ------------------------------------------------------------------ update warps, q <= 4: SIMT
// With at most 4 right-hand sides half of every update DMMA is padding (N = 8), and a warp issues a DMMA only every
// ~27 cycles: 128 useful FMAs per 27 issue cycles.  Plain DFMA does 32 FMAs per instruction: the 4 update warps then
// need ~1/3 of the issue time (and the FP64 pipe sees half the work), and stop being the role the tile waits for
// (ncu, profiles/r2_ug_ncu_s13.txt: update warps busy 85 % with the DMMA version).
//   NH = 2 (R <= 64):  lane = 16 h + (row mod 16): the two halves of a warp take the even / odd columns of the same 16
//                      rows (two 128-byte runs per load: conflict-free), partial sums meet by one shuffle;
//   NH = 1:            lane = row mod 32, RP = 1 (R <= 128) or 2 rows per lane (row, row + 128).
// B comes as warp-wide broadcasts from the zero-padded [column][8] copy, so all 4 accumulators are always computed.
__device__ __forceinline__ void lds_b4(uint32_t addr, double (&b)[4]) {
  asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(b[0]), "=d"(b[1]) : "r"(addr));
  asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(b[2]), "=d"(b[3]) : "r"(addr + 16u));
}

template <int NH, int RP>
__device__ __forceinline__ void update_role_simt(const UGArgs& u, const UGSmem& g, UGBars& bar, unsigned char* xbuf0,
                                                 unsigned char* tile0, const double* Bs, int64_t t_begin,
                                                 int ntile_cta, int warp, int lane, double (*dred)[16]) {
  const int R = u.rows, q = u.q, nchunk = u.nchunk;
  const uint32_t colstride = (uint32_t)R * 8u;
  const int h = NH == 2 ? lane >> 4 : 0;
  int rowt[RP];
  rowt[0] = NH == 2 ? 16 * warp + (lane & 15) : 32 * warp + lane;
  if (RP == 2) rowt[RP - 1] = rowt[0] + 128;
  uint32_t a_off[RP];
#pragma unroll
  for (int i = 0; i < RP; ++i) a_off[i] = (uint32_t)min(rowt[i], R - 1) * 8u + (uint32_t)h * colstride;  // (rows >= R: clamped, not stored)
  const uint32_t b_addr = smem_u32(Bs) + (uint32_t)h * 64u;
  double dself[4] = {0.0, 0.0, 0.0, 0.0}, dold[4] = {0.0, 0.0, 0.0, 0.0};
  int tb = 0;
  uint32_t par = 0;
  for (int t = 0; t < ntile_cta; ++t) {
    const int64_t r0 = (t_begin + t) * R;
    double acc[RP][4];
#pragma unroll
    for (int i = 0; i < RP; ++i)
#pragma unroll
      for (int c = 0; c < 4; ++c) acc[i][c] = 0.0;
    const uint32_t tile_addr = smem_u32(tile0) + (uint32_t)tb * g.tile_bytes;
    for (int cch = 0; cch < nchunk; ++cch) {
      mbar_wait(&bar.full[tb][cch], par);
      UG_TRACE(warp, lane, 10 + cch, t);
#ifndef B200_UG_NO_UPDATE
      const uint32_t ca = tile_addr + (uint32_t)(kCW * cch) * colstride, cb = b_addr + (uint32_t)(kCW * cch) * 64u;
      constexpr int CPC = kCW / NH;  // columns of a box per lane
      double a[CPC][RP];
#pragma unroll
      for (int jj = 0; jj < CPC; ++jj)
#pragma unroll
        for (int i = 0; i < RP; ++i) a[jj][i] = lds_f64(ca + (uint32_t)(NH * jj) * colstride + a_off[i]);
#pragma unroll
      for (int jj = 0; jj < CPC; ++jj) {
        double b[4];
        lds_b4(cb + (uint32_t)(NH * jj) * 64u, b);
#pragma unroll
        for (int i = 0; i < RP; ++i)
#pragma unroll
          for (int c = 0; c < 4; ++c) acc[i][c] = fma(a[jj][i], b[c], acc[i][c]);
      }
#endif
      if (((cch + 1) & (kGC - 1)) == 0 && cch + 1 < nchunk) {  // a release group handed back (the last one: with the X buffer)
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar.gempty[tb][(cch + 1) / kGC - 1]);
      }
    }
    if (NH == 2) {
#pragma unroll
      for (int c = 0; c < 4; ++c) acc[0][c] += __shfl_xor_sync(0xffffffffu, acc[0][c], 16);
    }
    // The old X tile came with the last box (rows >= n zero-filled); X_new replaces it in place for the Gram warps.
    double* xb = reinterpret_cast<double*>(xbuf0 + (size_t)tb * g.x_bytes);
#pragma unroll
    for (int i = 0; i < RP; ++i) {
      const int rt = rowt[i];
      const int64_t row = r0 + rt;
      if (rt < R && h == 0) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          if (c < q) {
            const double s = acc[i][c];
            double o;
            if (u.beta == 0.0) {
              o = row < u.n ? u.alpha * s : 0.0;
              dself[c] = fma(o, o, dself[c]);
            } else {
              const double v = xb[c * R + rt];
              o = fma(u.alpha, s, u.beta * v);
              dself[c] = fma(o, o, dself[c]);
              dold[c] = fma(o, v, dold[c]);
            }
            if (row < u.n) u.V[(size_t)c * u.ldv + row] = o;
            xb[c * R + rt] = o;
          }
        }
      }
    }
    __syncwarp();
    UG_TRACE(warp, lane, 2, t);
    if (lane == 0) {
      mbar_arrive(&bar.xfull[tb]);
      mbar_arrive(&bar.gempty[tb][(nchunk - 1) / kGC]);
    }
    if (++tb == u.nbuf) { tb = 0; par ^= 1u; }
  }
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const double a = warp_sum(dself[c]), b = warp_sum(dold[c]);
    if (lane == 0) {
      dred[warp][2 * c] = a;
      dred[warp][2 * c + 1] = b;
    }
  }
  if (lane < 8) dred[warp][8 + lane] = 0.0;  // (columns 4..7 do not exist on this path)
  // the arrivals that match the Gram warps' reduction rounds (gram_role)
#pragma unroll
  for (int round = 0; round <= kNG / 4; ++round) consumer_bar<(kNU + kNG) * 32>();
}

// ------------------------------------------------------------------ Gram warps
// NCB = 8-column blocks of A (compile time: no predicated slots, see update_role).  Warp wg takes the 4-row steps
// g = wg (mod 8) of every tile for all blocks.
// Few blocks: two accumulator sets alternate by step, so that consecutive DMMAs of a block do not wait for each other.
// one 4-row step (row = its byte offset in a column) against the blocks [C0, C0 + N) of A: N fragment loads, N DMMAs
template <int C0, int N, int NACC>
__device__ __forceinline__ void gram_step(double (&acc)[NACC][2], uint32_t a_addr, uint32_t x_addr, uint32_t row,
                                          uint32_t blk_stride, bool has_rhs) {
  const double b = has_rhs ? lds_f64(x_addr + row) : 0.0;
  double a[N];
#pragma unroll
  for (int cb = 0; cb < N; ++cb) a[cb] = lds_f64(a_addr + row + (uint32_t)(C0 + cb) * blk_stride);
#pragma unroll
  for (int cb = 0; cb < N; ++cb) B200_DMMA(acc[C0 + cb][0], acc[C0 + cb][1], a[cb], b);
}

// the sweep of one tile over release group GRP (8 blocks = kGC boxes), then the groups after it
template <int NCB, int GRP>
__device__ __forceinline__ void gram_groups(double (&gacc)[NCB][2], double (&acc1)[NCB <= 4 ? NCB : 1][2], UGBars& bar, int tb,
                                            uint32_t a_addr, uint32_t x_addr, uint32_t blk_stride, bool has_rhs, int ng,
                                            int wg, int lane) {
  constexpr int C0 = 8 * GRP, N = NCB - C0 < 8 ? NCB - C0 : 8, NS = NCB <= 4 ? 2 : 1;
  int g4 = wg;
  for (; g4 + kNG * (NS - 1) < ng; g4 += kNG * NS) {
    gram_step<C0, N>(gacc, a_addr, x_addr, (uint32_t)g4 * 32u, blk_stride, has_rhs);
    if constexpr (NS == 2) gram_step<C0, N>(acc1, a_addr, x_addr, (uint32_t)(g4 + kNG) * 32u, blk_stride, has_rhs);
  }
  if (NS == 2 && g4 < ng) gram_step<C0, N>(gacc, a_addr, x_addr, (uint32_t)g4 * 32u, blk_stride, has_rhs);  // odd step left over
  __syncwarp();
  if (lane == 0) mbar_arrive(&bar.gempty[tb][GRP]);  // (the last group also releases the tile's X buffer)
  if constexpr (C0 + 8 < NCB) gram_groups<NCB, GRP + 1>(gacc, acc1, bar, tb, a_addr, x_addr, blk_stride, has_rhs, ng, wg, lane);
}

// NCB = 8-column blocks of A (compile time: no predicated slots, see update_role).  Warp wg takes the 4-row steps
// g = wg (mod 8) of every tile, release group by release group (64 columns: 8 independent accumulator chains), so the
// producer refills a buffer while the sweep is still on its later columns.
// Few blocks: two accumulator sets alternate by step, so that consecutive DMMAs of a block do not wait for each other.
template <int NCB>
__device__ __forceinline__ void gram_role(const UGArgs& u, const UGSmem& g, UGBars& bar, unsigned char* xbuf0,
                                          unsigned char* tile0, int ntile_cta, int wg, int lane) {
  constexpr int NS = NCB <= 4 ? 2 : 1;
  const int R = u.rows;
  const int ng = R >> 2;
  const uint32_t colstride = (uint32_t)R * 8u;
  // A^T fragment: lane l = (basis column l/4 of the block, row l%4 of the step); V_new: (k = l%4, n = l/4)
  const uint32_t frag = (uint32_t)(lane >> 2) * colstride + (uint32_t)(lane & 3) * 8u;
  const bool has_rhs = (lane >> 2) < u.q;
  const uint32_t blk_stride = 8u * colstride;
  double gacc[NCB][2], acc1[NS == 2 ? NCB : 1][2];
#pragma unroll
  for (int cb = 0; cb < NCB; ++cb) gacc[cb][0] = gacc[cb][1] = 0.0;
#pragma unroll
  for (int cb = 0; cb < (NS == 2 ? NCB : 1); ++cb) acc1[cb][0] = acc1[cb][1] = 0.0;
  int tb = 0;
  uint32_t par = 0;
  for (int t = 0; t < ntile_cta; ++t) {
    // X_new complete: every update warp has waited for every box of the tile before it arrived here, so the boxes
    // are complete and (acquire through this barrier) visible
    mbar_wait(&bar.xfull[tb], par);
    UG_TRACE(kNU + wg, lane, 3, t);
    const uint32_t a_addr = smem_u32(tile0) + (uint32_t)tb * g.tile_bytes + frag;
    const uint32_t x_addr = smem_u32(xbuf0) + (uint32_t)tb * g.x_bytes + frag;
#ifndef B200_UG_NO_GRAM
    gram_groups<NCB, 0>(gacc, acc1, bar, tb, a_addr, x_addr, blk_stride, has_rhs, ng, wg, lane);
    UG_TRACE(kNU + wg, lane, 4, t);
#else
    __syncwarp();
    if (lane == 0) for (int grp = 0; grp <= (u.nchunk - 1) / kGC; ++grp) mbar_arrive(&bar.gempty[tb][grp]);
#endif
    if (++tb == u.nbuf) { tb = 0; par ^= 1u; }
  }
  if (NS == 2) {
#pragma unroll
    for (int cb = 0; cb < NCB; ++cb) {
      gacc[cb][0] += acc1[cb][0];
      gacc[cb][1] += acc1[cb][1];
    }
  }
  // CTA partial of A^T V_new: the 8 Gram warps' blocks are summed in shared memory in two rounds of four warps (the
  // tile buffers are free once every role has passed the first barrier; update_role makes the matching arrivals)
  // C fragment: lane l holds (basis column 8 cb + l/4, right-hand sides 2 (l%4) + {0,1})
  double* red = reinterpret_cast<double*>(tile0);  // [warp mod 4][8 NCB columns][8]
  consumer_bar<(kNU + kNG) * 32>();
#pragma unroll
  for (int round = 0; round < kNG / 4; ++round) {
    if ((wg >> 2) == round) {
#pragma unroll
      for (int cb = 0; cb < NCB; ++cb) {
        double* r = red + ((size_t)((wg & 3) * NCB + cb) * 8 + (lane >> 2)) * 8 + 2 * (lane & 3);
        r[0] = round == 0 ? gacc[cb][0] : r[0] + gacc[cb][0];
        r[1] = round == 0 ? gacc[cb][1] : r[1] + gacc[cb][1];
      }
    }
    consumer_bar<(kNU + kNG) * 32>();
  }
}

// run-time block count -> compile-time instantiation
template <int LO, int HI>
__device__ __forceinline__ void gram_dispatch(int ncb, const UGArgs& u, const UGSmem& g, UGBars& bar,
                                              unsigned char* xbuf0, unsigned char* tile0, int ntile_cta, int wg,
                                              int lane) {
  if constexpr (LO == HI) {
    gram_role<LO>(u, g, bar, xbuf0, tile0, ntile_cta, wg, lane);
  } else {
    constexpr int MID = (LO + HI) / 2;
    if (ncb <= MID) gram_dispatch<LO, MID>(ncb, u, g, bar, xbuf0, tile0, ntile_cta, wg, lane);
    else gram_dispatch<MID + 1, HI>(ncb, u, g, bar, xbuf0, tile0, ntile_cta, wg, lane);
  }
}

__global__ void __launch_bounds__(kThreadsUG, 1)
    update_gram_kernel(const __grid_constant__ UGArgs u, const __grid_constant__ CUtensorMap tmA,
                       const __grid_constant__ CUtensorMap tmV) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ UGBars bar;
  __shared__ double dred[kNU][16];

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int R = u.rows, p = u.p, q = u.q, nchunk = u.nchunk;
  const UGSmem g(R, nchunk, q);
  unsigned char* base = reinterpret_cast<unsigned char*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
  double* Bs = reinterpret_cast<double*>(base);
  unsigned char* xbuf0 = base + g.bs_bytes;
  unsigned char* tile0 = xbuf0 + (size_t)u.nbuf * g.x_bytes;
  const int64_t t_begin = u.ntiles * blockIdx.x / gridDim.x;
  const int64_t t_end = u.ntiles * (blockIdx.x + 1) / gridDim.x;
  const int ntile_cta = (int)(t_end - t_begin);

  for (int e = threadIdx.x; e < nchunk * kCW * 8; e += kThreadsUG) {
    const int j = e >> 3, c = e & 7;
    Bs[e] = (j < p && c < q) ? (u.B ? u.B[(size_t)c * u.ldb + j] : u.bval.v[c * p + j]) : 0.0;
  }
  if (threadIdx.x == 0) {
    for (int b = 0; b < u.nbuf; ++b) {
      for (int c = 0; c < nchunk; ++c) mbar_init(&bar.full[b][c], 1);
      for (int c = 0; c < kMaxChunks / kGC; ++c) mbar_init(&bar.gempty[b][c], kNU + kNG);
      mbar_init(&bar.xfull[b], kNU);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  if (warp == kNU + kNG) {
    // ------------------------------------------------------------------ producer (one thread)
    if (lane == 0) {
      const uint64_t pol = l2_policy_evict_first();
      const uint64_t pol_v = l2_policy_evict_last();  // X is the next kernel's operand
      int tb = 0;
      uint32_t par = 0;
      for (int t = 0; t < ntile_cta; ++t) {
        const int r0 = (int)((t_begin + t) * R);
        unsigned char* dst = tile0 + (size_t)tb * g.tile_bytes;
        for (int c = 0; c < nchunk; ++c) {
          if ((c & (kGC - 1)) == 0) {
            mbar_wait(&bar.gempty[tb][c / kGC], par ^ 1u);  // both roles are done with this group
            UG_TRACE(kNU + kNG, lane, 5 + c / kGC, t);
          }
          // the old X tile (beta != 0) travels with the LAST box (its buffer is released with the last group):
          // [column][R], the layout X_new is left in
          const bool with_x = c == nchunk - 1 && u.beta != 0.0;
          mbar_expect_tx(&bar.full[tb][c], g.chunk_bytes + (with_x ? (uint32_t)(q * R) * 8u : 0u));
          tma_load_2d(dst + (size_t)c * g.chunk_bytes, &tmA, r0, c * kCW, &bar.full[tb][c], pol);
          if (with_x) tma_load_2d(xbuf0 + (size_t)tb * g.x_bytes, &tmV, r0, 0, &bar.full[tb][c], pol_v);
        }
        if (++tb == u.nbuf) { tb = 0; par ^= 1u; }
      }
    }
    __syncwarp();
  } else if (warp < kNU && q <= 4 && !u.dmma_update) {
    if (R <= 64) update_role_simt<2, 1>(u, g, bar, xbuf0, tile0, Bs, t_begin, ntile_cta, warp, lane, dred);
    else if (R <= 128) update_role_simt<1, 1>(u, g, bar, xbuf0, tile0, Bs, t_begin, ntile_cta, warp, lane, dred);
    else update_role_simt<1, 2>(u, g, bar, xbuf0, tile0, Bs, t_begin, ntile_cta, warp, lane, dred);
  } else if (warp < kNU) {
    switch ((R + 4) >> 5) {  // row groups per update warp
      case 1: update_role<1>(u, g, bar, xbuf0, tile0, Bs, t_begin, ntile_cta, warp, lane, dred); break;
      case 2: update_role<2>(u, g, bar, xbuf0, tile0, Bs, t_begin, ntile_cta, warp, lane, dred); break;
      case 3: update_role<3>(u, g, bar, xbuf0, tile0, Bs, t_begin, ntile_cta, warp, lane, dred); break;
      case 4: update_role<4>(u, g, bar, xbuf0, tile0, Bs, t_begin, ntile_cta, warp, lane, dred); break;
      case 5: update_role<5>(u, g, bar, xbuf0, tile0, Bs, t_begin, ntile_cta, warp, lane, dred); break;
      case 6: update_role<6>(u, g, bar, xbuf0, tile0, Bs, t_begin, ntile_cta, warp, lane, dred); break;
      case 7: update_role<7>(u, g, bar, xbuf0, tile0, Bs, t_begin, ntile_cta, warp, lane, dred); break;
      default: update_role<8>(u, g, bar, xbuf0, tile0, Bs, t_begin, ntile_cta, warp, lane, dred); break;
    }
  } else {
    gram_dispatch<1, kMaxCB>((p + 7) >> 3, u, g, bar, xbuf0, tile0, ntile_cta, warp - kNU, lane);
  }
  __syncthreads();  // (the producer warp joins; the compute warps left their partials behind the named barrier)

  // ---- CTA partials (the roles have left them in shared memory: red = the tile buffers, dred) -> workspace
  const int ncb8 = (p + 7) >> 3;
  const double* red = reinterpret_cast<const double*>(tile0);
  const int ngram = p * q, nent = ngram + 2 * q;
  double* wsc = u.fd.ws + (size_t)blockIdx.x * nent;
  for (int e = threadIdx.x; e < nent; e += kThreadsUG) {
    double s = 0.0;
    if (e < ngram) {
      const int k = e / p, col = e - k * p;
#pragma unroll
      for (int w = 0; w < 4; ++w) s += red[((size_t)w * ncb8 * 8 + col) * 8 + k];
    } else {
#pragma unroll
      for (int w = 0; w < kNU; ++w) s += dred[w][e - ngram];
    }
    wsc[e] = s;
  }
  const bool fin = fold_partials_tall<kThreadsUG>(u.fd.ws, u.fd.ticket, nent, (int)gridDim.x, (int)blockIdx.x,
                                                  [&](int e, double s) {
                                                    if (e < ngram) u.fd.out[2 * q + e] = s;
                                                    else u.fd.out[e - ngram] = s;
                                                  });
  if (!fin) return;
  publish_result(u.fd.pub);
}

// rows per tile: R = 28 (mod 32); at least two tile buffers (+ the staged B + the X_new buffers) must fit.  Among
// the candidates that fit, the one that leaves the shortest idle tail when the tiles are dealt to the CTAs
// (larger tiles amortise the per-tile hand-overs: a 4 % longer tail is accepted for them).
int ug_bufs(int R, int p, int q) {
  const UGSmem g(R, (p + kCW - 1) / kCW, q);
  int nb = kMaxNB;
  while (nb >= 2 && g.total(nb) > kSmemUG) --nb;
  return nb >= 2 ? nb : 0;
}

int ug_pick_rows(int p, int q, int64_t n, int num_sms) {
  int best = 0;
  int64_t min_cost = 0;
  int64_t cost[8];
  int cand[8];
  int nc = 0;
  for (int R = 28; R <= 252; R += 32) {  // 32 K - 4: the 8-row groups deal evenly to the 4 update warps
    if (ug_bufs(R, p, q) == 0) break;
    const int64_t nt = (n + R - 1) / R;
    const int64_t grid = std::min<int64_t>(num_sms, nt);
    cand[nc] = R;
    cost[nc] = (nt + grid - 1) / grid * R;
    if (nc == 0 || cost[nc] < min_cost) min_cost = cost[nc];
    ++nc;
  }
  for (int i = 0; i < nc; ++i)
    if (cost[i] * 100 <= min_cost * 104) best = cand[i];
  return best;
}

}  // namespace

static int spec_mode() {  // B200_SPEC_GRAM: 0 never, 1 (default) when the call pattern predicts a hit, 2 always
  static const int m = [] {
    const char* e = getenv("B200_SPEC_GRAM");
    return e ? atoi(e) : 1;
  }();
  return m;
}

// can launch_update_gram take this shape at all?
bool update_gram_fits(Ctx* ctx, int p, int q, int64_t n) {
  if (spec_mode() == 0 || !ctx->fuse) return false;
  // (below ~24 columns the two plain kernels are as fast: their fixed costs are smaller, profiles/r2_ug_sweep_s12.txt)
  if (p < 24 || p > kMaxCB * 8 || q < 1 || q > 8) return false;
  if (n < (int64_t)ctx->num_sms * 28 * 4) return false;  // a few tiles per CTA, or the pipeline never fills
  return ug_pick_rows(p, q, n, ctx->num_sms) > 0;
}
bool update_gram_always() { return spec_mode() >= 2; }

// V = alpha A B + beta V as launch_update does; in addition fd.out[0 .. 2q) = {x_new.x_new, x_new.x_old}
// per column and fd.out[2q + k p + col] = sum_rows A(:, col) V_new(:, k)
int launch_update_gram(Ctx* ctx, double* V, int64_t ldv, int q, const double* A, int64_t lda, int p,
                       int64_t n, const double* dB, int ldb, double alpha, double beta,
                       const FusedDots& fd, const double* hB) {
  static thread_local UGArgs u;  // (holds the 8 KB parameter copy of B: not on the stack of every call)
  u.V = V; u.B = dB;
  if (hB) {
    if ((size_t)p * q > (size_t)kBParamDoubles) return fail(B200_ERR_ARG, "MvTimesMatAddMv: dense block too large for a kernel parameter (internal)");
    u.B = nullptr;
    std::memcpy(u.bval.v, hB, sizeof(double) * (size_t)p * q);
  } u.ldv = ldv; u.n = n; u.p = p; u.q = q; u.ldb = ldb;
  u.alpha = alpha; u.beta = beta; u.fd = fd;
  int R = ug_pick_rows(p, q, n, ctx->num_sms);
  if (const char* e = getenv("B200_UG_ROWS")) R = atoi(e);
  if (R <= 0) return fail(B200_ERR_ARG, "MvTimesMatAddMv: fused update + Gram does not fit (internal)");
  u.rows = R;
  u.nchunk = (p + kCW - 1) / kCW;
  u.nbuf = ug_bufs(R, p, q);
  // q <= 4: the SIMT update role up to 80 basis columns, the DMMA one beyond (measured cross-over, 1M rows: p = 64
  // 122 vs 142 us, p = 100 214 vs 204 us, p = 200 373 vs 336 us; profiles/r2_ug_roles_s19.txt).  B200_UG_DMMA_UPDATE=0/1 forces one.
  u.dmma_update = getenv("B200_UG_DMMA_UPDATE") ? atoi(getenv("B200_UG_DMMA_UPDATE")) : (p > 80 ? 1 : 0);
  if (const char* e = getenv("B200_UG_BUFS")) u.nbuf = std::max(2, std::min(u.nbuf, atoi(e)));
  if (u.nbuf < 2 || R % 32 != 28 || R > 252)
    return fail(B200_ERR_ARG, "MvTimesMatAddMv: fused update + Gram: bad tile height (internal)");
  u.ntiles = (n + R - 1) / R;
  const int grid = (int)std::max<int64_t>(1, std::min<int64_t>(ctx->num_sms, u.ntiles));
  const size_t nent = (size_t)p * q + 2 * (size_t)q;
  if (nent * (grid + grid / 16 + 1) * sizeof(double) > kWorkspaceBytes)
    return fail(B200_ERR_ARG, "MvTimesMatAddMv: partial-sum workspace too small");
  CUtensorMap tmA, tmV;
  B200_TRY(make_tmap(&tmA, A, n, p, lda, R, kCW));
  B200_TRY(make_tmap(&tmV, V, n, q, ldv, R, q));
  static unsigned attr_devs = 0;  // bit per device: the attribute is per device, not per process
  if (!(attr_devs & (1u << (ctx->device & 31)))) {
    B200_CUDA(cudaFuncSetAttribute(update_gram_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)kSmemUG));
    attr_devs |= 1u << (ctx->device & 31);
  }
  ctx->t_begin(K_UPDATE_GRAM, 8.0 * n * ((double)p + (beta == 0.0 ? 1.0 : 2.0) * q));
  update_gram_kernel<<<grid, kThreadsUG, UGSmem(R, u.nchunk, q).total(u.nbuf), ctx->stream>>>(u, tmA, tmV);
  ++ctx->launches;
  ctx->t_end();
  B200_CUDA(cudaGetLastError());
#ifdef B200_UG_TRACE
  if (const char* fn = getenv("B200_UG_TRACE_FILE")) {
    static unsigned long long h[13][4096][2];
    unsigned int hn[13];
    B200_CUDA(cudaStreamSynchronize(ctx->stream));
    B200_CUDA(cudaMemcpyFromSymbol(h, g_ug_trace, sizeof(h)));
    B200_CUDA(cudaMemcpyFromSymbol(hn, g_ug_trace_n, sizeof(hn)));
    unsigned int zero[13] = {};
    B200_CUDA(cudaMemcpyToSymbol(g_ug_trace_n, zero, sizeof(zero)));
    if (FILE* f = fopen(fn, "w")) {
      fprintf(f, "# p=%d q=%d R=%d nbuf=%d nchunk=%d\n", p, q, R, u.nbuf, u.nchunk);
      for (int w = 0; w < 13; ++w)
        for (unsigned int i = 0; i < hn[w] && i < 4096; ++i)
          fprintf(f, "%d %llu %d %d\n", w, h[w][i][0], (int)(h[w][i][1] >> 32), (int)(h[w][i][1] & 0xffffffffu));
      fclose(f);
    }
  }
#endif
  return 0;
}

}  // namespace b200
