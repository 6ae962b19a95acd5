// Device runtime of the B200 Block GMRES path: context (streams, memory pool,
// host mailbox, partial-sum workspace, NCCL), multivector and operator objects.
#ifndef B200_RUNTIME_CUH
#define B200_RUNTIME_CUH

#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <map>
#include <memory>
#include <vector>

#include "b200_belos.h"
#include "internal.h"

namespace b200 {

constexpr int kNumSM = 148;            // B200: 2 dies x 74 SMs
constexpr int kLdAlign = 32;           // leading dimension padded to 32 doubles (256 B)
constexpr size_t kMailboxDoubles = 1u << 16;   // 512 KB host-mapped result mailbox
constexpr size_t kWorkspaceBytes = 256u << 20; // per-CTA partial sums
constexpr size_t kTicketBytes = 256u << 10;    // global + per-group arrival counters (two-level folds)
constexpr int kStageSlots = 16;
// Fused-reduction bookkeeping: kernels that write a column (SpMM, MvTimesMatAddMv) also emit the
// column dot products Belos asks for next (x.x, x.x_old); the results wait in the upper part of
// the host mailbox until the matching MvDot / MvNorm call picks them up without a launch.
constexpr int kDotCache = 256;                  // remembered column dot products
constexpr int kAuxSlots = 32;                   // mailbox slots for results of fused epilogues
constexpr int kAuxSlotDoubles = 64;             // <= 32 columns x {self, with-old}
constexpr size_t kAuxBase = kMailboxDoubles - (size_t)kAuxSlots * kAuxSlotDoubles;
// Two larger slots below them for the fused update + Gram kernel (update_gram.cu): 2q dot products followed
// by the p x q block A^T V_new (p <= 256, q <= 8) that the next MvTransMv is expected to ask for.
constexpr int kSpecSlots = 2;
constexpr int kSpecSlotDoubles = 2112;
constexpr size_t kSpecBase = kAuxBase - (size_t)kSpecSlots * kSpecSlotDoubles;
constexpr size_t kStageSlotBytes = 512u << 10; // host B blocks for MvTimesMatAddMv
constexpr size_t kPackBytes = 1u << 20;        // the same block in DMMA fragment order (16 -> 24 columns)

#define B200_CUDA(call)                                                                   \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess)                                                                \
      return ::b200::failf(B200_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call,      \
                           cudaGetErrorString(e_));                                       \
  } while (0)

#define B200_TRY(call)            \
  do {                            \
    int rc_ = (call);             \
    if (rc_ != 0) return rc_;     \
  } while (0)

struct NcclApi;  // dlopen'ed function table (runtime.cu)

// peer-memory exchange windows (see Ctx::xw).  Window layout: [parity][source rank][kXchgMax] 16-byte
// LL words (value + sequence tag in one store, device_utils.cuh publish_result).
constexpr int kMaxPeers = 8;
constexpr int kXchgMax = 8192;   // entries per exchange; longer results travel in chunks
constexpr size_t kXchgBytes = (size_t)2 * kMaxPeers * kXchgMax * 16;
struct XchgWindow {
  void* win[kMaxPeers] = {};     // win[r]: rank r's window as mapped in this process
  int nranks = 1, rank = 0;
};

// How the finishing CTA of a reduction kernel hands its result to the host.  nranks <= 1: release
// `flag` with `seq` (results were written straight into the mailbox).  nranks > 1: run the
// peer-memory exchange of reduce.cu in this CTA (put src[0..count) into every rank's window, wait
// for the peers, fold in rank order into dst = mailbox) and then release the flag -- the separate
// exchange kernel and its launch gap disappear from the critical path of every reduction.
struct XchgPub {
  uint64_t* flag = nullptr;
  uint64_t seq = 0;
  int nranks = 1, rank = 0, count = 0, is_max = 0;
  uint64_t xseq = 0;
  const double* src = nullptr;
  double* dst = nullptr;
  uint64_t* err = nullptr;
  void* win[kMaxPeers] = {};
};

// a small dense operand handed to a kernel as a parameter (no staging copy, no H2D copy in front of the launch)
constexpr int kBParamDoubles = 1024;
struct BParam { double v[kBParamDoubles]; };

// where a kernel with a fused reduction epilogue puts its column dot products
struct FusedDots {
  double* out = nullptr;        // where the finishing CTA stores the sums (nullptr: no epilogue): the mailbox slot
                                // (single rank) or this rank's partial-sum slot (multi-rank)
  uint64_t seq = 0;
  double* ws = nullptr;         // per-CTA partials
  unsigned int* ticket = nullptr;
  // how the finishing CTA of the LAST launch of the operation hands the slot over: single rank -- release
  // the host flag with seq; multi-rank -- the peer exchange of publish_result() (all `count` entries of the
  // slot, rank-ordered fold into the mailbox) inside the producing kernel, so no exchange kernel is launched
  // behind it.  Launches that are not the last one carry a default (inert) XchgPub.
  XchgPub pub;
  void not_last() { pub = XchgPub(); }
  // multi-column update kernels (q <= 4): also emit new_i . new_k for i < k at out[2q + pair(i, k)]
  int cross = 0;
};

enum KClass { K_SPMM = 0, K_GRAM = 1, K_UPDATE = 2, K_REDUCE = 3, K_AXPBY = 4, K_HALO = 5, K_UPDATE_GRAM = 6, K_FBSTEP = 7 };

struct Ctx {
  int device = 0;
  int num_sms = kNumSM;               // queried at context creation
  cudaStream_t stream = nullptr;       // compute stream (all kernels)
  cudaStream_t comm_stream = nullptr;  // halo exchange
  cudaEvent_t ev_comm = nullptr, ev_compute = nullptr;
  // size-bucketed caching pool: Clone/CloneCopy are called several times per Arnoldi step
  std::multimap<size_t, void*> pool;
  size_t pool_bytes = 0;
  // host-mapped mailbox for reduction results + completion flag
  double* mbox_h = nullptr;
  double* mbox_d = nullptr;
  volatile uint64_t* flag_h = nullptr;
  uint64_t* flag_d = nullptr;
  uint64_t seq = 0;
  // device scratch
  double* ws = nullptr;            // per-CTA partials
  unsigned int* ticket = nullptr;  // last-block election counter (self-resetting)
  double* red_d = nullptr;         // device-side result (multi-rank: allreduce in place)
  double* aux_d = nullptr;         // multi-rank: per-rank partials of fused epilogues (kAuxSlots slots)
  double* dB = nullptr;            // device copy of the host dense block B
  double* dBpack = nullptr;        // B re-laid in DMMA fragment order (update.cu)
  char* stage_h = nullptr;         // pinned staging ring for B
  cudaEvent_t stage_ev[kStageSlots] = {};
  bool stage_used[kStageSlots] = {};
  int stage_next = 0;
  void* flush_buf = nullptr;       // > L2 scratch for bench hygiene
  // Peer-memory exchange for the small allreduces (multi-rank, one node): every rank's window is
  // mapped into every other rank through CUDA IPC; a one-CTA kernel stores this rank's partial
  // result into all windows over NVLink, raises a flag per peer, waits for the peers' flags and
  // folds the contributions in rank order (bitwise identical on every rank) straight into the
  // host mailbox.  Replaces ncclAllReduce + publish kernel (~30 us) by one ~6 us kernel.
  XchgWindow xw;
  void* xwin = nullptr;            // this rank's window (cudaMalloc, IPC-exported)
  uint64_t xseq = 0;               // exchange counter (parity selects the window half; low 32 bits = LL tag)
  bool p2p = false;
  // NCCL
  NcclApi* nccl = nullptr;
  void* comm = nullptr;
  int rank = 0, nranks = 1;
  // Park-Miller state for MvRandom
  unsigned int seed = 1;
  // content ids + remembered dot products (see kDotCache)
  uint64_t next_id = 1;
  struct DotEntry {
    uint64_t a = 0, b = 0;    // content ids (a <= b)
    double val = 0.0;
    uint64_t seq = 0;         // 0: val is final; else: mul * mbox_h[src], published once flag >= seq
    size_t src = 0;
    double mul = 1.0;         // (a column that was scaled after the product was launched: deferred MvScale)
  };
  DotEntry dots[kDotCache];
  int dot_next = 0;
  int aux_next = 0;
  bool fuse = true;           // B200_NO_FUSE=1 disables the fused epilogues and the dot memo
  int64_t dot_hits = 0;
  int64_t pending = 0;        // deferred scales + lazy copies outstanding (0: the prep_* hooks return at once)
  int64_t lazy_saved = 0, scale_folded = 0;  // launches that never happened (diagnostics)
  // Gram memo: A^T V of the listed columns was produced by the fused update + Gram kernel and waits at
  // mbox_h[src + k p + col] (once flag >= seq); consulted by MvTransMv.
  struct GramMemo {
    std::vector<uint64_t> a_ids, v_ids;
    uint64_t seq = 0;
    size_t src = 0;
    bool valid = false;
  } gmemo;
  // Call-pattern predictor for that kernel: did the MvTransMv(A, V) that the last eligible update could have
  // served actually come?  spec_next[h] = outcome that followed outcome history h (two calls deep) last time.
  struct SpecState {
    std::vector<uint64_t> a_ids, v_ids;  // the last eligible update (whether it speculated or not)
    bool open = false;                   // ... and no MvTransMv has matched it yet
    unsigned hist = 0;
    bool next_hit[4] = {true, true, true, true};
  } spec;
  int spec_next = 0;                     // next big slot
  int64_t gram_hits = 0, gram_spec = 0;  // memo answers / speculative launches (diagnostics)
  void remember_dot(uint64_t a, uint64_t b, double val, uint64_t seq = 0, size_t src = 0, double mul = 1.0);
  // column `from` was scaled by `factor` and is now known as `to`: its remembered products follow
  void rescale_dots(uint64_t from, uint64_t to, double factor);
  // the block the last multi-column MvTimesMatAddMv wrote (findBasis works through its columns next)
  // (identified by the storage's serial number, never by its address: a freed storage's address can be handed out
  // again on one rank and not on another, and the ranks must take the same decisions)
  struct LastBlock {
    uint64_t uid = 0;
    int c0 = 0, q = 0;
  } last_block;
  uint64_t next_uid = 1;
  int64_t skinny_hits = 0;   // one-column MvTransMv calls answered from remembered products
  double answered_bytes = 0.0;  // SURVEY 8(d) bytes of the calls that were answered without a kernel (memos)
  // true (and *val set, waiting for the producing kernel if needed) if the product is known
  bool recall_dot(uint64_t a, uint64_t b, double* val, int* rc);
  // counters / timing
  int64_t launches = 0, syncs = 0;
  bool timing = false;
  struct Stamp { cudaEvent_t a, b; int cls; };
  std::vector<Stamp> stamps;
  std::vector<cudaEvent_t> ev_free;
  double t_ms[B200_NUM_KCLASS] = {};
  int64_t t_calls[B200_NUM_KCLASS] = {};
  double t_bytes[B200_NUM_KCLASS] = {};

  int alloc(size_t bytes, void** out);
  void release(void* p, size_t bytes);
  // begin/end a timed region for kernel class cls (no-op unless timing enabled)
  void t_begin(int cls, double bytes);
  void t_end();
  int t_drain();
  // wait until the device has published sequence number `s` into the flag
  int wait_flag(uint64_t s);
  // multi-rank: allreduce red_d[0..count) (sum or max) then publish to mailbox; single-rank:
  // the kernel already wrote the mailbox.  Returns after the host can read mbox_h.
  int finish_reduction(int count, bool is_max);
  // announce the size of the reduction whose kernels are launched next (after ++seq), so that the
  // last of them can run the peer exchange itself (make_pub(true))
  void plan_reduction(int count, bool is_max);
  XchgPub make_pub(bool last);
  int xplan_count = 0;
  bool xplan_max = false, xplan_ok = false, xplan_done = false;
  double* result_ptr() { return nranks > 1 ? red_d : mbox_d; }
  uint64_t* result_flag() { return nranks > 1 ? nullptr : flag_d; }
};

struct Storage {
  Ctx* ctx;
  double* ptr;
  size_t bytes;
  int64_t ld = 0, n_local = 0;
  // content id per storage column: equal ids <=> bitwise equal columns.  Every kernel that
  // writes a column gives it a fresh id; pure copies propagate the source id.
  std::vector<uint64_t> ver;
  bool external = false;  // a raw device pointer was handed out: contents are no longer tracked
  uint64_t uid = 0;       // serial number of this storage within its context
  // Launch-bound single-column work of DGKS findBasis (BelosDGKSOrthoManager.hpp:1037, :1217-1222), kept
  // off the stream until something needs the data (capi.cu: prep_read / prep_write):
  //  * deferred MvScale of one column: its logical content is pscale[c] x the stored data; the skinny Gram
  //    that reads the column next folds the factor in and writes the scaled column back, any other reader
  //    applies it with a scale launch first;
  //  * lazy CloneCopy of one column: column c of this storage is a copy of src:col that was not made
  //    yet.  The copy that findBasis takes of column j is only ever read by one MvDot, which the dot memo
  //    answers; the one-column update that overwrites the source stores the old values into the copy
  //    on its way (no launch), and a copy that is released unread is never made.
  struct Lazy {
    std::shared_ptr<Storage> src;
    int col = -1;
  };
  std::vector<double> pscale;
  std::vector<Lazy> lazy;
  std::vector<std::vector<std::pair<Storage*, int>>> deps;  // pending lazy copies OF column c
  Storage(Ctx* c, double* p, size_t b, int ncols, int64_t ld_, int64_t n_)
      : ctx(c), ptr(p), bytes(b), ld(ld_), n_local(n_), ver((size_t)std::max(ncols, 1)),
        pscale((size_t)std::max(ncols, 1), 1.0), lazy((size_t)std::max(ncols, 1)), deps((size_t)std::max(ncols, 1)) {
    for (auto& v : ver) v = c->next_id++;
    uid = c->next_uid++;
  }
  ~Storage();
  double* colptr(int c) const { return ptr + (size_t)c * ld; }
};

struct MV {
  Ctx* ctx;
  std::shared_ptr<Storage> st;
  int64_t n_local, n_global, row_offset, ld;
  std::vector<int> cols;  // column index (inside the storage) of each vector of this object
  int ncols() const { return (int)cols.size(); }
  double* col(int j) const { return st->ptr + (size_t)cols[j] * ld; }
  const int* colv() const { return cols.data(); }
  // content id of column j (a fresh one each time if the storage is not tracked)
  uint64_t id(int j) const { return st->external ? ctx->next_id++ : st->ver[cols[j]]; }
  void touch(int j, int len = 1) const {
    for (int i = 0; i < len; ++i) st->ver[cols[j + i]] = ctx->next_id++;
  }
  void copy_id(int j, const MV* src, int js, int len = 1) const {
    uint64_t tmp[64];
    for (int i0 = 0; i0 < len; i0 += 64) {  // two phases: src and dst may be the same storage
      const int m = std::min(64, len - i0);
      for (int i = 0; i < m; ++i) tmp[i] = src->id(js + i0 + i);
      for (int i = 0; i < m; ++i) st->ver[cols[j + i0 + i]] = tmp[i];
    }
  }
};

// maximal run [j, j+len) on which every listed multivector advances by exactly one
// storage column per step (so the run is a plain strided block in each of them)
inline int run_length(int j, int n, const MV* a, int ja, const MV* b = nullptr, int jb = 0,
                      const MV* c = nullptr, int jc = 0) {
  int len = 1;
  while (j + len < n) {
    if (a && a->cols[ja + len] != a->cols[ja + len - 1] + 1) break;
    if (b && b->cols[jb + len] != b->cols[jb + len - 1] + 1) break;
    if (c && c->cols[jc + len] != c->cols[jc + len - 1] + 1) break;
    ++len;
  }
  return len;
}

struct HaloPeer {
  int rank;
  int64_t send_off, send_cnt;  // rows of mine the peer needs (offset into send_idx)
  int64_t recv_off, recv_cnt;  // ghost rows I get from the peer (offset into ghost tail)
  int64_t remote_off = 0;      // where my rows start in the PEER's ghost tail
  int64_t remote_nghost = 0;   // the peer's ghost count (leading dimension of its ghost block)
};

// Peer-memory halo ("put") arguments.  Every rank owns a halo window
//   [256 B: one uint64 flag per source rank][ghost block, parity 0][ghost block, parity 1]
// mapped into its neighbours through CUDA IPC.  The put kernel gathers the rows a neighbour needs
// and stores them straight into that neighbour's ghost block over NVLink (final [rhs][ghost]
// layout), then raises the neighbour's flag; the boundary-row kernel spins on its own flags.
struct HaloPut {
  int npeers = 0;
  int64_t send_off[kMaxPeers], send_cnt[kMaxPeers];
  double* dst[kMaxPeers];       // peer's ghost block (current parity) + remote_off
  int64_t dst_ld[kMaxPeers];    // peer's n_ghost
  uint64_t* flag[kMaxPeers];    // peer's flag word for this rank
};
struct HaloWait {
  int n = 0;
  const uint64_t* flag[kMaxPeers];
  uint64_t seq = 0;
  uint64_t* err = nullptr;   // host-visible error word (set when a neighbour never arrives)
};

// ---- window-staged SpMM (spmm_win.cu): per-tile descriptor and per-operator plan --------------
struct WinTile {
  int64_t blob_off;        // byte offset of the tile's blob in WinPlan::blob_d (128-B aligned)
  int32_t row0, nrows;     // first local row, rows (<= R)
  int32_t width;           // stored entries per row (sliced-ELL width of this tile)
  int32_t nruns, wlen;     // x-window runs and their total length (elements per right-hand side)
  int32_t ghost_mask;      // bit j: run j lies in the ghost block (run_start relative to it)
  int32_t run_start[8];    // first column of the run (multiple of 16)
  int32_t run_len[8];      // elements (multiple of 16)
  int32_t run_off[8];      // offset of the run inside the window
};
struct WinPlan {
  int R = 0;               // rows per tile (0: not built yet)
  bool ok = false;         // false: the format does not apply to this matrix
  int64_t ntiles = 0;
  WinTile* tiles_d = nullptr;
  unsigned char* blob_d = nullptr;
  size_t blob_bytes = 0;
  int32_t* int_tiles_d = nullptr;  // multi-rank: tiles without / with ghost runs
  int32_t* bnd_tiles_d = nullptr;
  int32_t* mix_tiles_d = nullptr;  // the interior tiles followed by the boundary tiles
  WinTile* tiles_mix_d = nullptr;  // the descriptors in that order (what the merged launch reads)
  int64_t n_int_tiles = 0, n_bnd_tiles = 0;
  int blob_cap = 0;        // largest tile blob (bytes, multiple of 128)
  int wcap = 0;            // largest window (elements, multiple of 16)
  double pad_ratio = 1.0;  // stored entries / nonzeros
};

// leading dimension of the ghost block: padded so that every right-hand side starts 256-B aligned
inline int64_t ghost_pad(int64_t n_ghost) { return (n_ghost + 31) & ~(int64_t)31; }

struct Op {
  Ctx* ctx;
  int64_t n_global, row_begin, row_end, n_local, n_ghost, nnz;
  int32_t* rowptr = nullptr;  // n_local+1
  int32_t* colind = nullptr;  // local column ids (owned first, then ghosts)
  double* values = nullptr;
  // interior / boundary split (boundary rows reference at least one ghost column)
  int32_t* bnd_rows = nullptr;
  int64_t n_bnd = 0;
  int32_t* int_rows = nullptr;  // only when n_ghost > 0
  int64_t n_int = 0;
  // halo
  std::vector<HaloPeer> peers;
  // true when ANY rank of the communicator has a halo neighbour: the apply path (and with it the
  // collectives it runs) is chosen from this globally agreed flag, never from the per-rank peer list --
  // a rank whose rows happen to be decoupled still walks through the same collective steps
  bool any_halo = false;
  int32_t* send_idx = nullptr;  // local row ids to pack, grouped by peer
  int64_t n_send = 0;
  double* send_buf = nullptr;   // n_send x k  (layout [rhs][row] per peer)
  double* xcol = nullptr;       // (n_local + n_ghost) x k column-map copy of x (multi-rank)
  int64_t xcol_ld = 0;
  int xcol_k = 0;
  double* recv_buf = nullptr;   // ghost block the kernels read (p2p: the current parity's half)
  size_t buf_k = 0;
  // peer-memory halo (see HaloPut)
  bool p2p_halo = false;
  void* hwin = nullptr;                 // own window
  void* peer_hwin[kMaxPeers] = {};      // peers[i]'s window as mapped here
  uint64_t hseq = 0;                    // applies done through the window (parity = hseq & 1)
  int64_t ghost_ld = 0;                 // ghost_pad(n_ghost)
  // host copy of the local CSR (rows sorted by local column id) for the window plans, built lazily
  std::vector<int32_t> h_rowptr, h_colind;
  std::vector<double> h_values;
  WinPlan win[2];                       // R = 256, 128
  int max_row_nnz = 0;
  // largest nonzero count of any tile of 64 / 128 / 256 consecutive rows (SpMM stage sizing)
  int64_t max_tile_nnz[3] = {0, 0, 0};
};

// ---- kernel launchers (defined in the .cu files) ---------------------------------
// elementwise.cu
int launch_fill(Ctx*, double* y, int64_t ld, int64_t n, int k, double a);
// y(:,j) = sthis*y(:,j) + sab * d(:) .* b(:,j)   (d: one column)
int launch_ewise_mul(Ctx*, double* y, int64_t ldy, const double* d, const double* b, int64_t ldb,
                     int64_t n, int k, double sab, double sthis);
// Chebyshev smoother passes (Ifpack_Chebyshev::ApplyInverse arithmetic)
int launch_cheby_first(Ctx*, double* w, int64_t ldw, double* y, int64_t ldy, const double* invd,
                       const double* x, int64_t ldx, int64_t n, int k, double oot, bool single);
int launch_cheby_step(Ctx*, double* w, int64_t ldw, double* y, int64_t ldy, const double* invd,
                      const double* x, int64_t ldx, const double* v, int64_t ldv, int64_t n, int k,
                      double d1, double d2);
struct Op;
int launch_diag(Ctx*, const Op* op, double* out, bool invert, double min_diag);
int launch_axpby(Ctx*, double* y, int64_t ldy, const double* a, int64_t lda, const double* b,
                 int64_t ldb, int64_t n, int k, double alpha, double beta);
int launch_scale_cols(Ctx*, double* y, int64_t ld, int64_t n, int k, const double* alpha_host);
int launch_random(Ctx*, double* y, int64_t ld, int64_t n, int k, int64_t row_offset,
                  int64_t n_global, int64_t seq_start, unsigned int seed);
int launch_l2_sweep(Ctx*, const double* buf, size_t bytes, double* sink);
// reduce.cu : out[j] for j<k written to ctx->result_ptr()+out_off
int launch_dot(Ctx*, const double* a, int64_t lda, const double* b, int64_t ldb, int64_t n, int k,
               int out_off, bool last);
int launch_norm(Ctx*, const double* a, int64_t lda, int64_t n, int k, int type, int out_off,
                bool last);
// out[j] = sum a(:,j)^2 into a device buffer, nothing published (per-rank partial of a fused slot)
int launch_dot_into(Ctx*, const double* a, int64_t lda, int64_t n, int k, double* out);
// gram.cu : C(p x q, ld=ldc) = alpha * A^T V into result_ptr()+out_off
// scales (skinny path only, see gram_is_skinny): p + 1 pending column factors (A's columns, then V's) that the
// kernel applies and writes back; nullptr: none
int launch_gram(Ctx*, const double* A, int64_t lda, int p, const double* V, int64_t ldv, int q,
                int64_t n, double alpha, int out_off, int ldc, bool last, const double* scales = nullptr);
bool gram_is_skinny(int p, int q);
// update.cu : V = alpha*A*B + beta*V with B (p x q, ld = p) already in ctx->dB + boff
// fused != nullptr: also emit {sum new^2, sum new*old} per column into fused->out[2k], [2k+1]
// save_old (q == 1, beta != 0 only): the kernel also stores V's old values there (a pending lazy copy of V)
// hB != nullptr (q <= 4 only): B is NOT in device memory; the compact host copy (p x q, ld = p) travels as a kernel
// parameter (p * q <= kBParamDoubles)
int launch_update(Ctx*, double* V, int64_t ldv, int q, const double* A, int64_t lda, int p,
                  int64_t n, const double* dB, int ldb, double alpha, double beta,
                  const FusedDots* fused = nullptr, double* save_old = nullptr, const double* hB = nullptr);
// findbasis.cu : one-column update with pending scales, lazy copy, and dots with the later columns of the block
bool fb_step_enabled();
int launch_fb_step(Ctx*, double* x, double* A, int64_t lda, int p, const double* coef, const double* scale,
                   const double* next, int64_t ldn, int nx, bool prev0_next, int64_t n, double alpha, double beta,
                   double* save_old, const FusedDots& fd);
// update_gram.cu : the same update; fused.out[0..2q) = the dot products, fused.out[2q + k p + col] = A^T V_new
bool update_gram_fits(Ctx*, int p, int q, int64_t n);
bool update_gram_always();
// hB != nullptr: B (compact, ld = p, p * q <= kBParamDoubles) from the host as a kernel parameter instead of dB
int launch_update_gram(Ctx*, double* V, int64_t ldv, int q, const double* A, int64_t lda, int p,
                       int64_t n, const double* dB, int ldb, double alpha, double beta,
                       const FusedDots& fused, const double* hB = nullptr);
// spmm.cu
// fused != nullptr (single rank, stream kernel only): also emit sum y(:,k)^2 into fused->out[k]
int launch_spmm(Ctx*, const Op* op, const int32_t* rows, int64_t nrows, const double* x,
                int64_t ldx, double* y, int64_t ldy, int k, const FusedDots* fused = nullptr);
int launch_pack(Ctx*, cudaStream_t s, const int32_t* idx, int64_t cnt, const double* x,
                int64_t ldx, int k, double* buf);
int launch_halo_put(Ctx*, const HaloPut& h, const int32_t* send_idx, int64_t n_send,
                    const double* x, int64_t ldx, int k, uint64_t seq);
// boundary rows after a peer-memory halo: the kernel waits for the neighbours' flags itself
int launch_spmm_boundary(Ctx*, const Op* op, const HaloWait& w, const double* x, int64_t ldx,
                         double* y, int64_t ldy, int k);

// spmm_win.cu
const WinPlan* win_plan_for(Ctx*, Op* op, int k, int* rc);
void free_win_plan(WinPlan* plan);
int spmm_win_grid(Ctx*, const WinPlan& pl, int k, int64_t ntiles);
int launch_spmm_win(Ctx*, const Op* op, const WinPlan& pl, const int32_t* tile_list, int64_t ntiles,
                    const double* x, int64_t ldx, double* y, int64_t ldy, int k, int k0, const FusedDots* fused,
                    int pstride, int part_base, int part_total, const HaloWait* wait, int* grid_out,
                    int64_t wait_from = 0);
// out[0..count) = 0 and hand over as a finished reduction (empty local row block)
int launch_zero_result(Ctx*, double* out, int count, const XchgPub& pub);

// the mailbox publish kernel used after a multi-rank allreduce
int launch_publish(Ctx*, int count);
// multi-rank with peer windows: exchange + fold + publish in one kernel (count <= kXchgMax)
int launch_xchg_publish(Ctx*, const double* src_d, double* dst, int count, bool is_max,
                        uint64_t seq);
// the same for an exchange that already has its numbers (XchgPub built by fused_slot / make_pub)
int launch_xchg_pub(Ctx*, const XchgPub& pub);

}  // namespace b200

struct b200_ctx : b200::Ctx {};
struct b200_mv : b200::MV {};
struct b200_op : b200::Op {};

#endif
