// Operator apply, window-staged variant: Y = A X with the x operand staged in shared memory.
//
// Reference semantics as spmm.cu: Epetra_CrsMatrix::Multiply(false,X,Y) -> GeneralMV / GeneralMM
// (epetra/src/Epetra_CrsMatrix.cpp:3315-3413, :3486-3585; Fortran strip-mined multi-RHS variant
// epetra/src/Epetra_dcrsmm.f:340-500): per row, sum of val*x[col] accumulated left to right.
//
// Why a second kernel.  The gather kernel of spmm.cu reads x through L1/L2 with 8-byte loads.  ncu
// (profiles/r1_spmm_tma_ncu_details_s18.txt) shows neither DRAM (41 %) nor L2 (29 %) nor the issue
// slots busy: 43 % of the gathers miss L1 by construction (a 7-point row touches three grid planes,
// only one of which the CTA's neighbouring rows share), and the number of L1 misses an SM can keep in
// flight bounds the kernel at ~62 % of the HBM roofline.  Here the x values a tile of R rows needs are
// moved by the bulk-copy engine instead (cp.async.bulk, the same async path that streams the matrix),
// which has no such limit:
//   * set-up (host, once per operator and tile height): the rows are cut into tiles of R rows; for each
//     tile the distinct columns are covered by at most 8 contiguous RUNS of x (16-element granules,
//     gaps < 64 elements merged; ghost columns form runs of their own in the halo block); the tile's
//     matrix entries are re-laid as a sliced-ELL blob [header | values[j][row] | window index[j][row]
//     (uint16) | row length[row] (uint8)] -- 10 bytes per stored entry instead of CSR's 12, one
//     contiguous range per tile;
//   * kernel: persistent CTAs own contiguous ranges of tiles.  One producer warp issues, per tile, one
//     bulk copy for the blob and one per (run, right-hand side) for the x window into an mbarrier
//     ring, several tiles ahead; consumer thread i reduces row i of the tile for all K right-hand
//     sides out of shared memory, left to right (the reference's order), conflict-free (consecutive
//     lanes read consecutive words of every array).
// Matrices whose tiles need more than 8 runs, more than 65535 window entries, rows longer than 255
// entries or more than 25 % padding keep the gather kernel (launch_spmm_win returns 1 = not applicable).
//
// Multi-rank: tiles without ghost runs ("interior") are launched next to the peer-memory halo put;
// tiles with ghost runs ("boundary") form a second launch of the same kernel that first waits for the
// neighbours' flags -- no row is computed twice.  The fused y.y epilogue (the MvDot(X,X) DGKS takes
// right after the apply, BelosDGKSOrthoManager.hpp:1269,1395) spans both launches.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "device_utils.cuh"
#include "runtime.cuh"

namespace b200 {

namespace {

constexpr int kMaxRuns = 8;
constexpr int kGran = 16;            // run granule (elements): 128-byte bulk copies
constexpr int kGapMerge = 4;         // granules: smaller gaps between runs are copied along
constexpr int kMaxStagesWin = 4;
constexpr size_t kWinSmemBudget = 200u << 10;  // per SM, all resident CTAs
constexpr int kHdrBytes = 128;

struct BlobHeader {  // first 128 bytes of a tile blob
  int32_t row0, nrows, width, pad;
};

// ------------------------------------------------------------------------------------------ set-up
struct TileScratch {
  std::vector<uint8_t> mark_own, mark_gh;  // granule bitmaps
  std::vector<int32_t> touched_own, touched_gh;
};

bool build_tile(const std::vector<int32_t>& rp, const std::vector<int32_t>& ci, int64_t n_local,
                int64_t row0, int R, TileScratch& sc, WinTile* td) {
  const int64_t r1 = std::min<int64_t>(row0 + R, n_local);
  td->row0 = (int32_t)row0;
  td->nrows = (int32_t)(r1 - row0);
  int width = 0;
  sc.touched_own.clear();
  sc.touched_gh.clear();
  for (int64_t r = row0; r < r1; ++r) {
    width = std::max(width, rp[r + 1] - rp[r]);
    for (int32_t e = rp[r]; e < rp[r + 1]; ++e) {
      const int32_t c = ci[e];
      if (c < n_local) {
        const int32_t g = c / kGran;
        if (!sc.mark_own[g]) { sc.mark_own[g] = 1; sc.touched_own.push_back(g); }
      } else {
        const int32_t g = (int32_t)(c - n_local) / kGran;
        if (!sc.mark_gh[g]) { sc.mark_gh[g] = 1; sc.touched_gh.push_back(g); }
      }
    }
  }
  for (int32_t g : sc.touched_own) sc.mark_own[g] = 0;
  for (int32_t g : sc.touched_gh) sc.mark_gh[g] = 0;
  td->width = width;
  td->nruns = 0;
  td->wlen = 0;
  td->ghost_mask = 0;
  if (width > 255) return false;
  auto add_runs = [&](std::vector<int32_t>& gs, bool ghost) {
    std::sort(gs.begin(), gs.end());
    size_t i = 0;
    while (i < gs.size()) {
      size_t j = i;
      while (j + 1 < gs.size() && gs[j + 1] - gs[j] <= kGapMerge) ++j;
      if (td->nruns >= kMaxRuns) return false;
      const int k = td->nruns++;
      td->run_start[k] = gs[i] * kGran;
      td->run_len[k] = (gs[j] - gs[i] + 1) * kGran;
      td->run_off[k] = td->wlen;
      td->wlen += td->run_len[k];
      if (ghost) td->ghost_mask |= 1 << k;
      i = j + 1;
    }
    return true;
  };
  if (!add_runs(sc.touched_own, false)) return false;
  if (!add_runs(sc.touched_gh, true)) return false;
  for (int k = td->nruns; k < kMaxRuns; ++k) td->run_start[k] = td->run_len[k] = td->run_off[k] = 0;
  return td->wlen <= 65535;
}

inline size_t blob_bytes_of(int width, int R) {
  return ((size_t)kHdrBytes + (size_t)width * R * 10 + R + 127) & ~(size_t)127;
}

void fill_blob(const std::vector<int32_t>& rp, const std::vector<int32_t>& ci, const std::vector<double>& va,
               int64_t n_local, int R, const WinTile& td, unsigned char* blob) {
  const size_t bytes = blob_bytes_of(td.width, R);
  std::memset(blob, 0, bytes);
  BlobHeader h{td.row0, td.nrows, td.width, 0};
  std::memcpy(blob, &h, sizeof(h));
  double* vals = reinterpret_cast<double*>(blob + kHdrBytes);
  uint16_t* lcol = reinterpret_cast<uint16_t*>(blob + kHdrBytes + (size_t)td.width * R * 8);
  uint8_t* cnt = blob + kHdrBytes + (size_t)td.width * R * 10;
  for (int i = 0; i < td.nrows; ++i) {
    const int64_t r = td.row0 + i;
    cnt[i] = (uint8_t)(rp[r + 1] - rp[r]);
    for (int32_t e = rp[r]; e < rp[r + 1]; ++e) {
      const int j = e - rp[r];
      const int32_t c = ci[e];
      const bool gh = c >= n_local;
      const int32_t cc = gh ? (int32_t)(c - n_local) : c;
      int slot = -1;
      for (int k = 0; k < td.nruns; ++k)
        if ((((td.ghost_mask >> k) & 1) != 0) == gh && cc >= td.run_start[k] && cc < td.run_start[k] + td.run_len[k]) {
          slot = td.run_off[k] + (cc - td.run_start[k]);
          break;
        }
      vals[(size_t)j * R + i] = va[e];
      lcol[(size_t)j * R + i] = (uint16_t)slot;
    }
  }
}

}  // namespace

// Host part of the plan: tiles + blob for tiles of R rows of a local CSR block (columns >= n_local are
// ghosts).  Returns false if the format does not apply (see the file header).
struct HostPlan {
  std::vector<WinTile> tiles;
  std::vector<unsigned char> blob;
  std::vector<int32_t> int_tiles, bnd_tiles;
  int blob_cap = 0, wcap = 0;
  double pad_ratio = 1.0;
};

static bool build_host_plan(const std::vector<int32_t>& rp, const std::vector<int32_t>& ci,
                            const std::vector<double>& va, int64_t n, int64_t n_ghost, int R, HostPlan* hp) {
  if (n <= 0 || rp.size() != (size_t)n + 1) return false;
  const int64_t ntiles = (n + R - 1) / R;
  const int64_t nnz = rp[(size_t)n];
  std::vector<WinTile>& tiles = hp->tiles;
  tiles.assign((size_t)ntiles, WinTile());
  const unsigned hw = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
  const unsigned nth = (unsigned)std::max<int64_t>(1, std::min<int64_t>(hw, ntiles / 64 + 1));
  std::vector<char> good(nth, 1);
  auto pass1 = [&](unsigned t) {
    TileScratch sc;
    sc.mark_own.assign((size_t)(n / kGran + 2), 0);
    sc.mark_gh.assign((size_t)(n_ghost / kGran + 2), 0);
    for (int64_t i = ntiles * t / nth; i < ntiles * (t + 1) / nth; ++i)
      if (!build_tile(rp, ci, n, i * R, R, sc, &tiles[(size_t)i])) {
        good[t] = 0;
        return;
      }
  };
  {
    std::vector<std::thread> th;
    for (unsigned t = 1; t < nth; ++t) th.emplace_back(pass1, t);
    pass1(0);
    for (auto& x : th) x.join();
  }
  for (char g : good)
    if (!g) return false;
  size_t off = 0, padded = 0;
  for (int64_t i = 0; i < ntiles; ++i) {
    WinTile& td = tiles[(size_t)i];
    td.blob_off = (int64_t)off;
    const size_t b = blob_bytes_of(td.width, R);
    off += b;
    padded += (size_t)td.width * td.nrows;
    hp->blob_cap = std::max(hp->blob_cap, (int)b);
    hp->wcap = std::max(hp->wcap, td.wlen);
    (td.ghost_mask ? hp->bnd_tiles : hp->int_tiles).push_back((int32_t)i);
  }
  hp->wcap = (hp->wcap + kGran - 1) / kGran * kGran;
  hp->pad_ratio = nnz > 0 ? (double)padded / (double)nnz : 1.0;
  if (hp->pad_ratio > 1.25) return false;
  hp->blob.assign(off + 256, 0);
  auto pass2 = [&](unsigned t) {
    for (int64_t i = ntiles * t / nth; i < ntiles * (t + 1) / nth; ++i)
      fill_blob(rp, ci, va, n, R, tiles[(size_t)i], hp->blob.data() + tiles[(size_t)i].blob_off);
  };
  {
    std::vector<std::thread> th;
    for (unsigned t = 1; t < nth; ++t) th.emplace_back(pass2, t);
    pass2(0);
    for (auto& x : th) x.join();
  }
  return true;
}

// Builds the window plan of `op` for tiles of R rows from the host copy of its CSR arrays
// (Op::h_rowptr / h_colind / h_values).  plan->ok == false: not applicable.
int build_win_plan(Op* op, int R, WinPlan* plan) {
  plan->R = R;
  plan->ok = false;
  HostPlan hp;
  if (!build_host_plan(op->h_rowptr, op->h_colind, op->h_values, op->n_local, op->n_ghost, R, &hp)) return 0;
  plan->pad_ratio = hp.pad_ratio;
  B200_CUDA(cudaMalloc((void**)&plan->tiles_d, hp.tiles.size() * sizeof(WinTile)));
  B200_CUDA(cudaMemcpy(plan->tiles_d, hp.tiles.data(), hp.tiles.size() * sizeof(WinTile), cudaMemcpyHostToDevice));
  B200_CUDA(cudaMalloc((void**)&plan->blob_d, hp.blob.size()));
  B200_CUDA(cudaMemcpy(plan->blob_d, hp.blob.data(), hp.blob.size(), cudaMemcpyHostToDevice));
  if (op->n_ghost > 0) {
    auto up = [&](const std::vector<int32_t>& v, int32_t** d) -> int {
      B200_CUDA(cudaMalloc((void**)d, std::max<size_t>(1, v.size()) * sizeof(int32_t)));
      if (!v.empty()) B200_CUDA(cudaMemcpy(*d, v.data(), v.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
      return 0;
    };
    B200_TRY(up(hp.int_tiles, &plan->int_tiles_d));
    B200_TRY(up(hp.bnd_tiles, &plan->bnd_tiles_d));
    std::vector<int32_t> mix(hp.int_tiles);
    mix.insert(mix.end(), hp.bnd_tiles.begin(), hp.bnd_tiles.end());
    B200_TRY(up(mix, &plan->mix_tiles_d));  // interior tiles, then boundary tiles: one launch (wait_from)
    // ... and the descriptors themselves in that order: the merged launch reads tiles_mix_d[t] directly.  (Through the
    // index list the producer warp had two dependent global loads per tile -- list entry, then descriptor -- on a
    // chain that is only one tile ahead: about the tile period itself (~1.6 us), and the multi-rank apply ran ~27 %
    // slower than the same rows on one rank: 156 vs 113 us at 4M rows, profiles/r2_apply_ab_n2_s34.txt.)
    std::vector<WinTile> tmix(mix.size());
    for (size_t i = 0; i < mix.size(); ++i) tmix[i] = hp.tiles[(size_t)mix[i]];
    B200_CUDA(cudaMalloc((void**)&plan->tiles_mix_d, std::max<size_t>(1, tmix.size()) * sizeof(WinTile)));
    if (!tmix.empty())
      B200_CUDA(cudaMemcpy(plan->tiles_mix_d, tmix.data(), tmix.size() * sizeof(WinTile), cudaMemcpyHostToDevice));
    plan->n_int_tiles = (int64_t)hp.int_tiles.size();
    plan->n_bnd_tiles = (int64_t)hp.bnd_tiles.size();
  }
  B200_CUDA(cudaDeviceSynchronize());
  plan->ntiles = (int64_t)hp.tiles.size();
  plan->blob_bytes = hp.blob.size();
  plan->blob_cap = hp.blob_cap;
  plan->wcap = hp.wcap;
  plan->ok = true;
  return 0;
}

void free_win_plan(WinPlan* plan) {
  cudaFree(plan->tiles_d);
  cudaFree(plan->blob_d);
  cudaFree(plan->int_tiles_d);
  cudaFree(plan->bnd_tiles_d);
  cudaFree(plan->mix_tiles_d);
  cudaFree(plan->tiles_mix_d);
  *plan = WinPlan();
}

}  // namespace b200

// Host-only view of the plan (tests): see include/b200_belos.h
extern "C" int b200_spmm_window_plan(int64_t n_local, int64_t n_ghost, const int32_t* rowptr,
                                     const int32_t* colind_local, const double* values, int rows_per_tile,
                                     int64_t* ntiles, int64_t* blob_bytes, int32_t* blob_cap, int32_t* wcap,
                                     void* tiles_out, void* blob_out) {
  using namespace b200;
  if (!rowptr || n_local < 0 || (rows_per_tile != 128 && rows_per_tile != 256))
    return fail(B200_ERR_ARG, "b200_spmm_window_plan: bad arguments");
  static_assert(sizeof(WinTile) == 128, "tile descriptor layout");
  const int64_t nnz = rowptr[n_local];
  std::vector<int32_t> rp(rowptr, rowptr + n_local + 1), ci(colind_local, colind_local + nnz);
  std::vector<double> va(values, values + nnz);
  HostPlan hp;
  if (!build_host_plan(rp, ci, va, n_local, n_ghost, rows_per_tile, &hp)) return 1;  // not applicable
  if (ntiles) *ntiles = (int64_t)hp.tiles.size();
  if (blob_bytes) *blob_bytes = (int64_t)hp.blob.size();
  if (blob_cap) *blob_cap = hp.blob_cap;
  if (wcap) *wcap = hp.wcap;
  if (tiles_out) std::memcpy(tiles_out, hp.tiles.data(), hp.tiles.size() * sizeof(WinTile));
  if (blob_out) std::memcpy(blob_out, hp.blob.data(), hp.blob.size());
  return 0;
}

namespace b200 {

namespace {

// ------------------------------------------------------------------------------------------ kernel
struct WinArgs {
  const WinTile* tiles;
  const int32_t* tile_list;   // nullptr: tiles 0 .. ntiles-1
  int64_t ntiles;             // entries of the list handled by this launch
  const unsigned char* blob;
  const double* x;            // owned columns
  const double* xg;           // ghost block (columns >= n_owned), leading dimension ldg
  double* y;
  int64_t ldx, ldg, ldy;
  int blob_cap, wcap, nstage;
  // fused y.y epilogue: per-CTA partials at ws[k * pstride + part_base + cta]; the launch with
  // part_total > 0 folds ws[k * pstride + 0 .. part_total) and publishes
  FusedDots fd;
  int pstride, part_base, part_total;
  HaloWait wait;              // neighbours' flags to wait for ...
  int64_t wait_from;          // ... before the first list entry >= wait_from (0: before anything; the entries from
                              // there on are the tiles with ghost runs)
};

__device__ __forceinline__ uint32_t lds_u16(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint32_t lds_u8(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}

template <int K, int R, bool FUSE>
__global__ void __launch_bounds__(R + 32)
    spmm_win_kernel(const __grid_constant__ WinArgs a) {
  constexpr int NWC = R / 32;  // consumer warps
  constexpr int UJ = K >= 8 ? 1 : 2;  // entries per row processed together
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint64_t full_bar[kMaxStagesWin], empty_bar[kMaxStagesWin];
  __shared__ double red[NWC][K];
  __shared__ int give_up;
  const uint32_t stage0 = (smem_u32(smem_raw) + 127u) & ~127u;
  unsigned char* const stage0p = smem_raw + (stage0 - smem_u32(smem_raw));
  const uint32_t stage_bytes = (uint32_t)a.blob_cap + (uint32_t)K * (uint32_t)a.wcap * 8u;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // Tiles are dealt ROUND-ROBIN: at any time the CTAs of the grid work inside one band of gridDim.x consecutive
  // tiles that sweeps through the rows, so the three x planes a 7-point tile reads are fetched from DRAM once and
  // served from L2 to the tiles of the neighbouring planes a few microseconds later.  (Contiguous tile ranges per
  // CTA put the re-use a whole kernel duration apart: at 8M rows x no longer fits L2 and was read from DRAM three
  // times -- 0.65 of the HBM roofline with the DRAM pipe saturated.)
  const int64_t t_begin = blockIdx.x, t_end = a.ntiles, t_step = gridDim.x;

  if (threadIdx.x == 0) {
    give_up = 0;
    for (int s = 0; s < a.nstage; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], NWC);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (a.wait.n > 0 && a.wait_from <= 0) {
    // peer-memory halo: the ghost block is complete once every neighbour's flag shows this apply
    if ((int)threadIdx.x < a.wait.n && !spin_acquire_sys(a.wait.flag[threadIdx.x], a.wait.seq, a.wait.err))
      give_up = 1;
    __syncthreads();
    if (give_up) return;  // a neighbour never arrived: store nothing, the host sees the error word
  }

  double ss[K];
#pragma unroll
  for (int k = 0; k < K; ++k) ss[k] = 0.0;

  if (warp == NWC) {
    // ---------------------------------------------------------------- producer warp
    // (generic-proxy acquire of the neighbours' stores above -> async-proxy reads below)
    asm volatile("fence.proxy.async;" ::: "memory");
    int s = 0;
    uint32_t ph = 0;
    const uint64_t pol_m = l2_policy_evict_first();  // the matrix is streamed once per apply
    const uint64_t pol_x = l2_policy_evict_last();   // x is re-read by the tiles of the neighbouring grid planes
    // The descriptor of tile t + 1 is fetched while tile t's copies are issued and its ring slot is awaited:
    // a global-load round trip per tile on the producer's critical path left the consumers waiting for data.
    struct Desc { int32_t width, nruns, wlen, gmask, rs, rl, ro; int64_t boff; };
    const int l8 = lane & 7;
    auto load_desc = [&](int64_t t) {
      Desc d;
      const int64_t tid = a.tile_list ? (int64_t)__ldg(a.tile_list + t) : t;
      const WinTile* td = a.tiles + tid;
      d.width = __ldg(&td->width); d.nruns = __ldg(&td->nruns); d.wlen = __ldg(&td->wlen);
      d.gmask = __ldg(&td->ghost_mask); d.boff = __ldg(&td->blob_off);
      d.rs = __ldg(&td->run_start[l8]); d.rl = __ldg(&td->run_len[l8]); d.ro = __ldg(&td->run_off[l8]);
      return d;
    };
    Desc nxt = load_desc(t_begin < t_end ? t_begin : 0);
    bool halo_ready = !(a.wait.n > 0 && a.wait_from > 0);
    for (int64_t t = t_begin; t < t_end; t += t_step) {
      if (!halo_ready && t >= a.wait_from) {
        // One launch for interior and boundary tiles: the list holds the tiles without ghost runs first, so the
        // neighbours' puts have the whole interior sweep to arrive; this warp waits for their flags only here, before
        // the first window copy out of the ghost block.  (A neighbour that never arrives: the error word is set, the
        // host reports B200_ERR_TIMEOUT; the boundary rows of this apply are then computed from stale ghosts.)
        if (lane < a.wait.n) spin_acquire_sys(a.wait.flag[lane], a.wait.seq, a.wait.err);
        __syncwarp();
        asm volatile("fence.proxy.async;" ::: "memory");
        halo_ready = true;
      }
      const Desc d = nxt;
      if (t + t_step < t_end) nxt = load_desc(t + t_step);
      const int32_t width = d.width, nruns = d.nruns, wlen = d.wlen, gmask = d.gmask, rs = d.rs, rl = d.rl, ro = d.ro;
      const int64_t boff = d.boff;
      const uint32_t bbytes = ((uint32_t)kHdrBytes + (uint32_t)width * R * 10u + R + 127u) & ~127u;
      mbar_wait(&empty_bar[s], ph ^ 1u);
      unsigned char* st = stage0p + (size_t)s * stage_bytes;
      if (lane == 0) {
        mbar_expect_tx(&full_bar[s], bbytes + (uint32_t)K * (uint32_t)wlen * 8u);
        bulk_load_1d(st, a.blob + boff, bbytes, &full_bar[s], pol_m);
      }
      __syncwarp();
      const int ncopies = nruns * K;
      for (int base = 0; base < ncopies; base += 32) {
        const int idx = base + lane;
        const int j = min(idx / K, kMaxRuns - 1), k = idx % K;
        const int32_t start = __shfl_sync(0xffffffffu, rs, j);
        const int32_t len = __shfl_sync(0xffffffffu, rl, j);
        const int32_t off = __shfl_sync(0xffffffffu, ro, j);
        if (idx < ncopies) {
          const double* src = ((gmask >> j) & 1) ? a.xg + (size_t)k * a.ldg + start : a.x + (size_t)k * a.ldx + start;
          bulk_load_1d(st + a.blob_cap + ((size_t)k * a.wcap + off) * 8, src, (uint32_t)len * 8u, &full_bar[s], pol_x);
        }
      }
      if (++s == a.nstage) { s = 0; ph ^= 1u; }
    }
    __syncwarp();
  } else {
    // ---------------------------------------------------------------- consumer warps: thread = row
    const int i = threadIdx.x;
    int s = 0;
    uint32_t ph = 0;
    for (int64_t t = t_begin; t < t_end; t += t_step) {
      mbar_wait(&full_bar[s], ph);
      const uint32_t st = stage0 + (uint32_t)s * stage_bytes;
      const int32_t row0 = lds_s32(st), nrows = lds_s32(st + 4), width = lds_s32(st + 8);
      const uint32_t sv = st + kHdrBytes + (uint32_t)i * 8u;
      const uint32_t sc = st + kHdrBytes + (uint32_t)width * (R * 8u) + (uint32_t)i * 2u;
      const int cnt = (int)lds_u8(st + kHdrBytes + (uint32_t)width * (R * 10u) + (uint32_t)i);
      const uint32_t sx = st + (uint32_t)a.blob_cap;
      const uint32_t kstride = (uint32_t)a.wcap * 8u;
      double acc[K];
#pragma unroll
      for (int k = 0; k < K; ++k) acc[k] = 0.0;
      // every lane walks the tile's full width (uniform trip count); entries past the row's own length
      // are padding: their window index is 0 (a staged, finite-or-not value) and their FMA is predicated off
      for (int j = 0; j < width; j += UJ) {
        double v[UJ];
        uint32_t c[UJ];
#pragma unroll
        for (int u = 0; u < UJ; ++u) {
          const bool in = j + u < width;
          v[u] = in ? lds_f64(sv + (uint32_t)(j + u) * (R * 8u)) : 0.0;
          c[u] = in ? lds_u16(sc + (uint32_t)(j + u) * (R * 2u)) : 0u;
        }
        double xv[UJ][K];
#pragma unroll
        for (int u = 0; u < UJ; ++u)
#pragma unroll
          for (int k = 0; k < K; ++k) xv[u][k] = lds_f64(sx + (uint32_t)k * kstride + c[u] * 8u);
#pragma unroll
        for (int u = 0; u < UJ; ++u)
          if (j + u < cnt) {
#pragma unroll
            for (int k = 0; k < K; ++k) acc[k] = fma(v[u], xv[u][k], acc[k]);
          }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty_bar[s]);
      if (++s == a.nstage) { s = 0; ph ^= 1u; }
      if (i < nrows) {
        double* yr = a.y + row0 + i;
#pragma unroll
        for (int k = 0; k < K; ++k) {
          yr[(size_t)k * a.ldy] = acc[k];
          if (FUSE) ss[k] = fma(acc[k], acc[k], ss[k]);
        }
      }
    }
  }
  if (!FUSE) return;
  if (warp < NWC) {
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const double t = warp_sum(ss[k]);
      if (lane == 0) red[warp][k] = t;
    }
  }
  __syncthreads();
  if (threadIdx.x < K) {
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < NWC; ++w) t += red[w][threadIdx.x];
    a.fd.ws[(size_t)threadIdx.x * a.pstride + a.part_base + blockIdx.x] = t;
  }
  if (a.part_total <= 0) return;  // an earlier launch of a split apply: the last one folds
  if (!last_cta_arrives(a.fd.ticket, gridDim.x)) return;
  // flat deterministic fold: warp k sums the partials of right-hand side k in a fixed order
  for (int k = warp; k < K; k += NWC + 1) {
    const double* p = a.fd.ws + (size_t)k * a.pstride;
    double t0 = 0.0, t1 = 0.0;
    int q = lane;
    for (; q + 32 < a.part_total; q += 64) {
      t0 += __ldcg(p + q);
      t1 += __ldcg(p + q + 32);
    }
    if (q < a.part_total) t0 += __ldcg(p + q);
    const double t = warp_sum(t0 + t1);
    if (lane == 0) a.fd.out[k] = t;
  }
  publish_result(a.fd.pub);
}

struct WinGeom { int ctas, nstage; size_t smem; };

// resident CTAs per SM and ring depth for K right-hand sides; ctas == 0: does not fit
WinGeom win_geom(const WinPlan& pl, int K) {
  WinGeom g{0, 0, 0};
  const size_t stage = (size_t)pl.blob_cap + (size_t)K * pl.wcap * 8;
  if (2 * stage + 128 > kWinSmemBudget) return g;
  int ctas = (int)std::min<size_t>(3, kWinSmemBudget / (2 * stage + 1280));
  ctas = std::min(ctas, 2048 / (pl.R + 32));
  static const int force_ctas = [] {
    const char* e = getenv("B200_WIN_CTAS");
    return e ? atoi(e) : 0;
  }();
  if (force_ctas > 0) ctas = std::min(ctas, force_ctas);
  if (ctas < 1 || ctas * pl.R < 256) return g;
  int nst = (int)((kWinSmemBudget / ctas - 1280) / stage);
  nst = std::max(2, std::min(nst, kMaxStagesWin));
  static const int force_st = [] {
    const char* e = getenv("B200_WIN_STAGES");
    return e ? atoi(e) : 0;
  }();
  if (force_st >= 2) nst = std::min(nst, force_st);
  g.ctas = ctas;
  g.nstage = nst;
  g.smem = (size_t)nst * stage + 128;
  return g;
}

template <int K, int R>
int win_launch_t(Ctx* ctx, const WinPlan& pl, const WinGeom& g, WinArgs& a, bool fuse, int grid) {
  static unsigned attr_dev_mask[2] = {0, 0};  // per device (bit) and per FUSE variant
  const unsigned bit = 1u << (ctx->device & 31);
  if (!(attr_dev_mask[fuse ? 1 : 0] & bit)) {
    if (fuse)
      B200_CUDA(cudaFuncSetAttribute(spmm_win_kernel<K, R, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)kWinSmemBudget + 128));
    else
      B200_CUDA(cudaFuncSetAttribute(spmm_win_kernel<K, R, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)kWinSmemBudget + 128));
    attr_dev_mask[fuse ? 1 : 0] |= bit;
  }
  if (fuse)
    spmm_win_kernel<K, R, true><<<grid, R + 32, g.smem, ctx->stream>>>(a);
  else
    spmm_win_kernel<K, R, false><<<grid, R + 32, g.smem, ctx->stream>>>(a);
  ++ctx->launches;
  B200_CUDA(cudaGetLastError());
  return 0;
}

template <int K>
int win_launch(Ctx* ctx, const WinPlan& pl, const WinGeom& g, WinArgs& a, bool fuse, int grid) {
  if (pl.R == 256) return win_launch_t<K, 256>(ctx, pl, g, a, fuse, grid);
  return win_launch_t<K, 128>(ctx, pl, g, a, fuse, grid);
}

}  // namespace

// Plan (built on first use) for k right-hand sides, or nullptr if the window format does not apply.
const WinPlan* win_plan_for(Ctx* ctx, Op* op, int k, int* rc) {
  *rc = 0;
  static const bool on = [] {
    const char* e = getenv("B200_SPMM_WIN");
    return !(e && e[0] == '0');
  }();
  if (!on || op->n_local <= 0 || (k != 1 && k != 2 && k != 4 && k != 8 && k != 16)) return nullptr;
  static const int force_r = [] {
    const char* e = getenv("B200_WIN_R");
    return e ? atoi(e) : 0;
  }();
  const int cand[2] = {256, 128};
  for (int c = 0; c < 2; ++c) {
    if (force_r && cand[c] != force_r) continue;
    WinPlan& pl = op->win[c];
    if (pl.R == 0) {
      if (op->h_rowptr.empty()) return nullptr;  // host arrays already dropped and this R never built
      *rc = build_win_plan(op, cand[c], &pl);
      if (*rc) return nullptr;
    }
    if (!pl.ok) continue;
    if (win_geom(pl, k).ctas > 0) return &pl;
  }
  return nullptr;
}

// One launch of the window kernel over `ntiles` entries of tile_list (nullptr: all tiles).
// (x, y: first of the k columns of this launch; k0: its index in the apply's block -- the ghost block holds all of them)
int launch_spmm_win(Ctx* ctx, const Op* op, const WinPlan& pl, const int32_t* tile_list, int64_t ntiles,
                    const double* x, int64_t ldx, double* y, int64_t ldy, int k, int k0, const FusedDots* fused,
                    int pstride, int part_base, int part_total, const HaloWait* wait, int* grid_out, int64_t wait_from) {
  const WinGeom g = win_geom(pl, k);
  if (g.ctas <= 0) return fail(B200_ERR_ARG, "window SpMM: geometry does not fit (internal)");
  const int grid = (int)std::max<int64_t>(1, std::min<int64_t>(ntiles, (int64_t)ctx->num_sms * g.ctas));
  if (grid_out) *grid_out = grid;
  if (ntiles <= 0) return 0;
  WinArgs a;
  a.tiles = pl.tiles_d;
  a.tile_list = tile_list;
  static const bool mix_by_list = [] {  // B200_SPMM_MIX_LIST=1: the merged launch goes through the index list (A/B)
    const char* e = getenv("B200_SPMM_MIX_LIST");
    return e && e[0] == '1';
  }();
  if (tile_list && tile_list == pl.mix_tiles_d && pl.tiles_mix_d && !mix_by_list) {
    a.tiles = pl.tiles_mix_d;  // the descriptors in list order: no indirection
    a.tile_list = nullptr;
  }
  a.ntiles = ntiles;
  a.blob = pl.blob_d;
  a.x = x;
  a.xg = op->recv_buf ? op->recv_buf + (size_t)k0 * op->ghost_ld : nullptr;
  a.y = y;
  a.ldx = ldx;
  a.ldg = op->ghost_ld;
  a.ldy = ldy;
  a.blob_cap = pl.blob_cap;
  a.wcap = pl.wcap;
  a.nstage = g.nstage;
  const bool fuse = fused && fused->out;
  if (fuse) a.fd = *fused;
  a.pstride = pstride;
  a.part_base = part_base;
  a.part_total = part_total;
  if (wait) a.wait = *wait;
  a.wait_from = wait_from;
  switch (k) {
    case 1: return win_launch<1>(ctx, pl, g, a, fuse, grid);
    case 2: return win_launch<2>(ctx, pl, g, a, fuse, grid);
    case 4: return win_launch<4>(ctx, pl, g, a, fuse, grid);
    case 8: return win_launch<8>(ctx, pl, g, a, fuse, grid);
    case 16: return win_launch<16>(ctx, pl, g, a, fuse, grid);
    default: return fail(B200_ERR_ARG, "window SpMM: unsupported block width (internal)");
  }
}

// number of CTAs launch_spmm_win would use for `ntiles` tiles (fused-epilogue bookkeeping)
int spmm_win_grid(Ctx* ctx, const WinPlan& pl, int k, int64_t ntiles) {
  const WinGeom g = win_geom(pl, k);
  return (int)std::max<int64_t>(1, std::min<int64_t>(ntiles, (int64_t)ctx->num_sms * std::max(1, g.ctas)));
}

}  // namespace b200
