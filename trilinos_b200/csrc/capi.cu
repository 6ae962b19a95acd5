// C-ABI entry points for multivectors and the CSR operator (include/b200_belos.h).
// Host-side logic only: argument checks with Epetra's error conventions, view/run
// bookkeeping, staging of the small dense blocks, halo choreography.  All arithmetic on the
// n-sized data happens in the kernels of elementwise.cu / reduce.cu / gram.cu / update.cu /
// spmm.cu; there is no host fallback.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <numeric>

#include "runtime.cuh"

namespace b200 {
int nccl_allreduce_i64(Ctx* ctx, int64_t* buf, size_t count, cudaStream_t s);
int nccl_group_start(Ctx* ctx);
int nccl_group_end(Ctx* ctx);
int nccl_send(Ctx* ctx, const void* buf, size_t count, bool f64, int peer, cudaStream_t s);
int nccl_recv(Ctx* ctx, void* buf, size_t count, bool f64, int peer, cudaStream_t s);
}  // namespace b200

using namespace b200;

namespace {

inline int64_t padded_ld(int64_t n) {
  int64_t ld = (std::max<int64_t>(n, 1) + kLdAlign - 1) / kLdAlign * kLdAlign;
  return ld;
}

int new_mv(Ctx* ctx, int64_t n_local, int64_t n_global, int64_t row_offset, int ncols,
           b200_mv** out) {
  if (ncols < 0 || n_local < 0) return fail(B200_ERR_ARG, "multivector: negative size");
  auto* m = new b200_mv();
  m->ctx = ctx;
  m->n_local = n_local;
  m->n_global = n_global;
  m->row_offset = row_offset;
  m->ld = padded_ld(n_local);
  const size_t bytes = (size_t)m->ld * std::max(ncols, 1) * sizeof(double);
  void* p = nullptr;
  int rc = ctx->alloc(bytes, &p);
  if (rc) {
    delete m;
    return rc;
  }
  m->st = std::make_shared<Storage>(ctx, (double*)p, bytes, ncols, m->ld, n_local);
  m->cols.resize(ncols);
  std::iota(m->cols.begin(), m->cols.end(), 0);
  *out = m;
  return 0;
}

inline bool same_map(const MV* a, const MV* b) {
  return a->ctx == b->ctx && a->n_local == b->n_local && a->n_global == b->n_global;
}

bool columns_overlap(const MV* a, const MV* b) {
  if (a->st.get() != b->st.get()) return false;
  for (int ca : a->cols)
    for (int cb : b->cols)
      if (ca == cb) return true;
  return false;
}

// A mailbox slot + sequence number for a kernel with a fused reduction epilogue of `count` results.  Single rank:
// the kernel's last CTA publishes straight into the mailbox.  Multi-rank (peer windows available): the kernel
// leaves this rank's partial sums in a device slot and its finishing CTA runs the peer exchange itself
// (publish_result), folding all ranks' partials into the mailbox -- no host round trip, no exchange kernel.
// The exchange numbers are taken here, i.e. in call order, which is the same on every rank.  Remembered
// products that were still waiting in the recycled slot are dropped.
bool fused_slot(Ctx* ctx, FusedDots* fd, size_t* src, int count, bool big = false) {
  if (!ctx->fuse || (ctx->nranks != 1 && !ctx->p2p)) return false;
  // big: one of the kSpecSlots larger slots below the small ones (fused update + Gram: dots, then A^T V_new)
  int slot;
  size_t base, cap, dev_off;
  if (big) {
    slot = ctx->spec_next;
    ctx->spec_next = (slot + 1) % kSpecSlots;
    base = kSpecBase + (size_t)slot * kSpecSlotDoubles;
    cap = kSpecSlotDoubles;
    dev_off = (size_t)kAuxSlots * kAuxSlotDoubles + (size_t)slot * kSpecSlotDoubles;
    if (ctx->gmemo.valid && ctx->gmemo.src >= base && ctx->gmemo.src < base + cap) ctx->gmemo.valid = false;
  } else {
    slot = ctx->aux_next;
    ctx->aux_next = (slot + 1) % kAuxSlots;
    base = kAuxBase + (size_t)slot * kAuxSlotDoubles;
    cap = kAuxSlotDoubles;
    dev_off = (size_t)slot * kAuxSlotDoubles;
  }
  for (int i = 0; i < kDotCache; ++i)
    if (ctx->dots[i].seq != 0 && ctx->dots[i].src >= base && ctx->dots[i].src < base + cap)
      ctx->dots[i] = Ctx::DotEntry();
  fd->seq = ++ctx->seq;
  fd->ws = ctx->ws;
  fd->ticket = ctx->ticket;
  fd->pub = XchgPub();
  fd->pub.flag = ctx->flag_d;
  fd->pub.seq = fd->seq;
  if (ctx->nranks == 1) {
    fd->out = ctx->mbox_d + base;
  } else {
    fd->out = ctx->aux_d + dev_off;
    fd->pub.nranks = ctx->nranks;
    fd->pub.rank = ctx->rank;
    fd->pub.count = count;
    fd->pub.xseq = ctx->xseq + 1;
    ctx->xseq += (uint64_t)std::max(1, (count + kXchgMax - 1) / kXchgMax);
    fd->pub.src = fd->out;
    fd->pub.dst = ctx->mbox_d + base;
    fd->pub.err = ctx->flag_d + 1;
    for (int i = 0; i < ctx->nranks; ++i) fd->pub.win[i] = ctx->xw.win[i];
  }
  *src = base;
  return true;
}

// For a fused slot whose sums were NOT produced by a kernel with the epilogue (a plain reduction launch filled the
// slot): multi-rank, queue the one-CTA exchange kernel with the numbers fused_slot took.
int fused_finish_separately(Ctx* ctx, const FusedDots& fd) {
  if (ctx->nranks == 1) return 0;
  return launch_xchg_pub(ctx, fd.pub);
}

// ---- deferred single-column scales and lazy single-column copies (see Storage in runtime.cuh) ------------------
// forget the pending-copy link of (st, c) without making the copy
void lazy_unlink(Storage* st, int c) {
  Storage::Lazy& lz = st->lazy[(size_t)c];
  if (!lz.src) return;
  auto& d = lz.src->deps[(size_t)lz.col];
  for (size_t i = 0; i < d.size(); ++i)
    if (d[i].first == st && d[i].second == c) {
      d.erase(d.begin() + (long)i);
      break;
    }
  lz.src.reset();
  lz.col = -1;
  --st->ctx->pending;
}

// make the copy now (the source still holds the data: whoever overwrites a source column calls
// deps_materialize first)
int lazy_materialize(Storage* st, int c) {
  Storage::Lazy& lz = st->lazy[(size_t)c];
  if (!lz.src) return 0;
  const Storage* s = lz.src.get();
  B200_TRY(launch_axpby(st->ctx, st->colptr(c), st->ld, s->colptr(lz.col), s->ld, nullptr, 0, st->n_local, 1, 1.0, 0.0));
  lazy_unlink(st, c);
  return 0;
}

int deps_materialize(Storage* st, int c) {
  while (!st->deps[(size_t)c].empty()) {
    const auto d = st->deps[(size_t)c].back();
    B200_TRY(lazy_materialize(d.first, d.second));
  }
  return 0;
}

int scale_flush(Storage* st, int c) {
  const double f = st->pscale[(size_t)c];
  if (f == 1.0) return 0;
  B200_TRY(launch_axpby(st->ctx, st->colptr(c), st->ld, st->colptr(c), st->ld, nullptr, 0, st->n_local, 1, f, 0.0));
  st->pscale[(size_t)c] = 1.0;
  --st->ctx->pending;
  return 0;
}

// Before a kernel READS columns [j0, j0 + len) of m: the data must be there.  keep_scale: the caller folds pending
// scale factors in itself (skinny Gram).
int prep_read(const MV* m, int j0 = 0, int len = -1, bool keep_scale = false) {
  if (m->ctx->pending == 0) return 0;
  if (len < 0) len = m->ncols() - j0;
  Storage* st = m->st.get();
  for (int j = j0; j < j0 + len; ++j) {
    const int c = m->cols[(size_t)j];
    B200_TRY(lazy_materialize(st, c));
    if (!keep_scale) B200_TRY(scale_flush(st, c));
  }
  return 0;
}

// Before a kernel WRITES columns [j0, j0 + len) of m.  overwrite: the kernel never reads their old contents.
int prep_write(const MV* m, int j0, int len, bool overwrite) {
  if (m->ctx->pending == 0) return 0;
  Storage* st = m->st.get();
  for (int j = j0; j < j0 + len; ++j) {
    const int c = m->cols[(size_t)j];
    B200_TRY(deps_materialize(st, c));
    if (overwrite) {
      lazy_unlink(st, c);
      if (st->pscale[(size_t)c] != 1.0) {
        st->pscale[(size_t)c] = 1.0;
        --st->ctx->pending;
      }
    } else {
      B200_TRY(lazy_materialize(st, c));
      B200_TRY(scale_flush(st, c));
    }
  }
  return 0;
}

// Widest block of right-hand sides the window SpMM kernel takes in one launch for a `len`-wide apply (0: none;
// *pl = its plan).  If the whole width does not fit the shared-memory window (16 right-hand sides of a 5-point
// stencil), half or a quarter of it may: the matrix is then streamed once per group, which for short rows (the
// matrix is a small share of the traffic) still beats the gather kernel (Recirc2D 2000^2: 0.85 vs 0.55 of the
// HBM roofline at 8 vs 16 right-hand sides, profiles/r2_spmm_tile_dealing_sweep_s3.txt).
int win_width(Ctx* ctx, Op* op, int len, const WinPlan** pl, int* rc) {
  *pl = win_plan_for(ctx, op, len, rc);
  if (*pl) return len;
  if (*rc || op->n_local <= 0 || (double)op->nnz / (double)op->n_local > 8.0) return 0;
  for (int kw = len / 2; kw >= 4; kw /= 2) {
    if (len % kw) continue;
    *pl = win_plan_for(ctx, op, kw, rc);
    if (*rc) return 0;
    if (*pl) return kw;
  }
  return 0;
}

// Park-Miller jump: seed * 16807^e mod (2^31-1)
unsigned int pm_advance(unsigned int seed, unsigned long long e) {
  unsigned long long r = seed % 2147483647ull, base = 16807ull;
  while (e) {
    if (e & 1ull) r = (r * base) % 2147483647ull;
    base = (base * base) % 2147483647ull;
    e >>= 1;
  }
  return (unsigned int)r;
}

}  // namespace

// =========================================================================== multivector
extern "C" int b200_mv_create(b200_ctx* ctx, int64_t n_local, int64_t n_global,
                              int64_t row_offset, int ncols, int zero_out, b200_mv** out) {
  if (!ctx || !out) return fail(B200_ERR_ARG, "b200_mv_create: null argument");
  if (n_global < n_local) return fail(B200_ERR_ARG, "b200_mv_create: n_global < n_local");
  B200_TRY(new_mv(ctx, n_local, n_global, row_offset, ncols, out));
  if (zero_out && ncols > 0)
    B200_CUDA(cudaMemsetAsync((*out)->st->ptr, 0, (size_t)(*out)->ld * ncols * sizeof(double),
                              ctx->stream));
  return 0;
}

extern "C" int b200_mv_destroy(b200_mv* mv) {
  delete mv;
  return 0;
}

extern "C" int b200_mv_clone(const b200_mv* mv, int ncols, b200_mv** out) {
  if (!mv || !out) return fail(B200_ERR_ARG, "b200_mv_clone: null argument");
  return new_mv(mv->ctx, mv->n_local, mv->n_global, mv->row_offset, ncols, out);
}

static int check_index(const b200_mv* mv, const int* index, int nindex, const char* who) {
  if (nindex < 0 || (nindex > 0 && !index)) return failf(B200_ERR_ARG, "%s: bad index list", who);
  for (int i = 0; i < nindex; ++i)
    if (index[i] < 0 || index[i] >= mv->ncols())
      return failf(B200_ERR_ARG, "%s: index[%d]=%d out of range [0,%d)", who, i, index[i],
                   mv->ncols());
  return 0;
}

extern "C" int b200_mv_clone_view(const b200_mv* mv, const int* index, int nindex,
                                  b200_mv** out) {
  if (!mv || !out) return fail(B200_ERR_ARG, "b200_mv_clone_view: null argument");
  B200_TRY(check_index(mv, index, nindex, "CloneView"));
  auto* v = new b200_mv();
  v->ctx = mv->ctx;
  v->st = mv->st;
  v->n_local = mv->n_local;
  v->n_global = mv->n_global;
  v->row_offset = mv->row_offset;
  v->ld = mv->ld;
  v->cols.resize(nindex);
  for (int i = 0; i < nindex; ++i) v->cols[i] = mv->cols[index[i]];
  *out = v;
  return 0;
}

extern "C" int b200_mv_clone_copy(const b200_mv* mv, const int* index, int nindex,
                                  b200_mv** out) {
  if (!mv || !out) return fail(B200_ERR_ARG, "b200_mv_clone_copy: null argument");
  std::vector<int> all;
  if (!index) {
    all.resize(mv->ncols());
    std::iota(all.begin(), all.end(), 0);
    index = all.data();
    nindex = (int)all.size();
  }
  B200_TRY(check_index(mv, index, nindex, "CloneCopy"));
  b200_mv* c = nullptr;
  B200_TRY(new_mv(mv->ctx, mv->n_local, mv->n_global, mv->row_offset, nindex, &c));
  if (nindex == 1 && mv->ctx->fuse && !mv->st->external && mv->n_local > 0) {
    // one column (DGKS findBasis' oldMXj, BelosDGKSOrthoManager.hpp:1037): the copy is made only if somebody
    // reads it or the source is about to change
    Storage* src = mv->st.get();
    const int sc = mv->cols[(size_t)index[0]];
    int rc = lazy_materialize(src, sc);   // (a copy of a pending copy: make the first one real)
    if (!rc) rc = scale_flush(src, sc);
    if (rc) {
      delete c;
      return rc;
    }
    c->st->lazy[0].src = mv->st;
    c->st->lazy[0].col = sc;
    src->deps[(size_t)sc].push_back({c->st.get(), 0});
    ++mv->ctx->pending;
    c->copy_id(0, mv, index[0]);
    *out = c;
    return 0;
  }
  {
    int rc = 0;
    for (int i = 0; i < nindex && !rc; ++i) rc = prep_read(mv, index[i], 1);
    if (rc) {
      delete c;
      return rc;
    }
  }
  // copy maximal runs of consecutive source columns
  int j = 0;
  while (j < nindex) {
    int len = 1;
    while (j + len < nindex && mv->cols[index[j + len]] == mv->cols[index[j + len - 1]] + 1) ++len;
    int rc = launch_axpby(mv->ctx, c->col(j), c->ld, mv->col(index[j]), mv->ld, nullptr, 0,
                          mv->n_local, len, 1.0, 0.0);
    if (rc) {
      delete c;
      return rc;
    }
    j += len;
  }
  for (int i = 0; i < nindex; ++i) c->copy_id(i, mv, index[i]);  // a copy has its source's content
  *out = c;
  return 0;
}

extern "C" int64_t b200_mv_global_length(const b200_mv* mv) { return mv ? mv->n_global : -1; }
extern "C" int64_t b200_mv_local_length(const b200_mv* mv) { return mv ? mv->n_local : -1; }
extern "C" int b200_mv_num_vecs(const b200_mv* mv) { return mv ? mv->ncols() : -1; }
extern "C" b200_ctx* b200_mv_ctx(const b200_mv* mv) {
  return mv ? static_cast<b200_ctx*>(mv->ctx) : nullptr;
}

extern "C" int b200_mv_init(b200_mv* mv, double alpha) {
  if (!mv) return fail(B200_ERR_ARG, "MvInit: null multivector");
  B200_TRY(prep_write(mv, 0, mv->ncols(), true));
  for (int j = 0; j < mv->ncols();) {
    const int len = run_length(j, mv->ncols(), mv, j);
    B200_TRY(launch_fill(mv->ctx, mv->col(j), mv->ld, mv->n_local, len, alpha));
    mv->touch(j, len);
    j += len;
  }
  return 0;
}

extern "C" int b200_mv_random(b200_mv* mv) {
  if (!mv) return fail(B200_ERR_ARG, "MvRandom: null multivector");
  Ctx* ctx = mv->ctx;
  B200_TRY(prep_write(mv, 0, mv->ncols(), true));
  for (int j = 0; j < mv->ncols();) {
    const int len = run_length(j, mv->ncols(), mv, j);
    B200_TRY(launch_random(ctx, mv->col(j), mv->ld, mv->n_local, len, mv->row_offset,
                           mv->n_global, (int64_t)j * mv->n_global, ctx->seed));
    mv->touch(j, len);
    j += len;
  }
  ctx->seed = pm_advance(ctx->seed, (unsigned long long)mv->n_global * mv->ncols());
  return 0;
}

extern "C" int b200_mv_add_mv(b200_mv* mv, double alpha, const b200_mv* A, double beta,
                              const b200_mv* B) {
  if (!mv || !A || !B) return fail(B200_ERR_ARG, "MvAddMv: null multivector");
  // Epetra_MultiVector::Update returns -1 on column mismatch, -2 on length mismatch
  if (A->ncols() != mv->ncols() || B->ncols() != mv->ncols())
    return fail(B200_ERR_ARG, "MvAddMv: number of vectors differs (Epetra Update: -1)");
  if (!same_map(mv, A) || !same_map(mv, B))
    return fail(B200_ERR_SHAPE, "MvAddMv: vector lengths differ (Epetra Update: -2)");
  Ctx* ctx = mv->ctx;
  const int k = mv->ncols();
  const int64_t n = mv->n_local;
  if (alpha != 0.0) B200_TRY(prep_read(A));
  if (beta != 0.0) B200_TRY(prep_read(B));
  B200_TRY(prep_write(mv, 0, k, false));
  // alpha == 0 / beta == 0 short-circuits must not read the dropped operand
  // (epetra/src/Epetra_MultiVector.cpp:1348-1355)
  for (int j = 0; j < k;) {
    const int len = run_length(j, k, mv, j, A, j, B, j);
    if (alpha == 0.0 && beta == 0.0) {
      B200_TRY(launch_fill(ctx, mv->col(j), mv->ld, n, len, 0.0));
      mv->touch(j, len);
    } else if (alpha == 0.0) {
      if (!(beta == 1.0 && B->col(j) == mv->col(j))) {
        B200_TRY(launch_axpby(ctx, mv->col(j), mv->ld, B->col(j), B->ld, nullptr, 0, n, len, beta,
                              0.0));
        if (beta == 1.0) mv->copy_id(j, B, j, len); else mv->touch(j, len);
      }
    } else if (beta == 0.0) {
      if (!(alpha == 1.0 && A->col(j) == mv->col(j))) {
        B200_TRY(launch_axpby(ctx, mv->col(j), mv->ld, A->col(j), A->ld, nullptr, 0, n, len,
                              alpha, 0.0));
        if (alpha == 1.0) mv->copy_id(j, A, j, len); else mv->touch(j, len);
      }
    } else {
      B200_TRY(launch_axpby(ctx, mv->col(j), mv->ld, A->col(j), A->ld, B->col(j), B->ld, n, len,
                            alpha, beta));
      mv->touch(j, len);
    }
    j += len;
  }
  return 0;
}

extern "C" int b200_mv_scale(b200_mv* mv, double alpha) {
  if (!mv) return fail(B200_ERR_ARG, "MvScale: null multivector");
  if (alpha == 1.0) return 0;
  if (mv->ncols() == 1 && mv->ctx->fuse && !mv->st->external && mv->n_local > 0) {
    // one column (the normalisation at the end of every findBasis column, BelosDGKSOrthoManager.hpp:1217-1222):
    // remember the factor; the skinny Gram that reads the column next applies it on the way, anything else
    // triggers the scale launch first
    Storage* st = mv->st.get();
    const int c = mv->cols[0];
    B200_TRY(prep_write(mv, 0, 1, false));  // (copies of the unscaled column are made now; an older factor is applied)
    st->pscale[(size_t)c] = alpha;
    ++mv->ctx->pending;
    const uint64_t before = mv->id(0);
    mv->touch(0, 1);
    // products with this column that are already known (or on their way) stay known: x . y -> alpha (x . y)
    mv->ctx->rescale_dots(before, mv->id(0), alpha);
    return 0;
  }
  B200_TRY(prep_write(mv, 0, mv->ncols(), false));
  for (int j = 0; j < mv->ncols();) {
    const int len = run_length(j, mv->ncols(), mv, j);
    B200_TRY(launch_axpby(mv->ctx, mv->col(j), mv->ld, mv->col(j), mv->ld, nullptr, 0,
                          mv->n_local, len, alpha, 0.0));
    mv->touch(j, len);
    j += len;
  }
  return 0;
}

extern "C" int b200_mv_scale_cols(b200_mv* mv, const double* alpha, int nalpha) {
  if (!mv || !alpha) return fail(B200_ERR_ARG, "MvScale: null argument");
  if (nalpha != mv->ncols())
    return failf(B200_ERR_ARG,
                 "MvScale: vector of scaling factors has length %d, multivector has %d columns",
                 nalpha, mv->ncols());
  B200_TRY(prep_write(mv, 0, mv->ncols(), false));
  for (int j = 0; j < mv->ncols();) {
    const int len = run_length(j, mv->ncols(), mv, j);
    B200_TRY(launch_scale_cols(mv->ctx, mv->col(j), mv->ld, mv->n_local, len, alpha + j));
    mv->touch(j, len);
    j += len;
  }
  return 0;
}

extern "C" int b200_mv_set_block(b200_mv* mv, const b200_mv* A, const int* index, int nindex) {
  if (!mv || !A) return fail(B200_ERR_ARG, "SetBlock: null multivector");
  B200_TRY(check_index(mv, index, nindex, "SetBlock"));
  if (A->ncols() < nindex) return fail(B200_ERR_ARG, "SetBlock: A has fewer columns than index");
  if (!same_map(mv, A)) return fail(B200_ERR_SHAPE, "SetBlock: vector lengths differ");
  B200_TRY(prep_read(A, 0, nindex));
  for (int i = 0; i < nindex; ++i) B200_TRY(prep_write(mv, index[i], 1, mv->col(index[i]) != A->col(i)));
  int j = 0;
  while (j < nindex) {
    int len = 1;
    while (j + len < nindex && mv->cols[index[j + len]] == mv->cols[index[j + len - 1]] + 1 &&
           A->cols[j + len] == A->cols[j + len - 1] + 1)
      ++len;
    if (mv->col(index[j]) != A->col(j))
      B200_TRY(launch_axpby(mv->ctx, mv->col(index[j]), mv->ld, A->col(j), A->ld, nullptr, 0,
                            mv->n_local, len, 1.0, 0.0));
    j += len;
  }
  for (int i = 0; i < nindex; ++i) mv->copy_id(index[i], A, i);
  return 0;
}

extern "C" int b200_mv_dot(const b200_mv* mv, const b200_mv* A, double* b, int nb) {
  if (!mv || !A || !b) return fail(B200_ERR_ARG, "MvDot: null argument");
  if (A->ncols() != mv->ncols()) return fail(B200_ERR_ARG, "MvDot: number of vectors differs");
  if (!same_map(mv, A)) return fail(B200_ERR_SHAPE, "MvDot: vector lengths differ");
  const int k = mv->ncols();
  if (nb < k) return 0;  // reference silently does nothing (BelosEpetraAdapter.cpp:505)
  if (k == 0) return 0;
  Ctx* ctx = mv->ctx;
  // products a fused epilogue (or an earlier call on unchanged columns) already produced
  std::vector<uint64_t> ia(k), ib(k);
  bool known = ctx->fuse;
  for (int j = 0; j < k; ++j) {
    ia[j] = mv->id(j);
    ib[j] = A->id(j);
  }
  for (int j = 0; j < k && known; ++j) {
    int rc = 0;
    known = ctx->recall_dot(ia[j], ib[j], &b[j], &rc);
    if (rc) return rc;
  }
  if (known) {
    ctx->dot_hits += k;
    ctx->answered_bytes += 8.0 * mv->n_local * k * (mv == A ? 1.0 : 2.0);
    return 0;
  }
  B200_TRY(prep_read(mv));
  B200_TRY(prep_read(A));
  ++ctx->seq;
  ctx->plan_reduction(k, false);
  for (int j = 0; j < k;) {
    const int len = run_length(j, k, mv, j, A, j);
    B200_TRY(launch_dot(ctx, mv->col(j), mv->ld, A->col(j), A->ld, mv->n_local, len, j,
                        j + len >= k));
    j += len;
  }
  B200_TRY(ctx->finish_reduction(k, false));
  for (int j = 0; j < k; ++j) {
    b[j] = ctx->mbox_h[j];
    ctx->remember_dot(ia[j], ib[j], b[j]);
  }
  return 0;
}

extern "C" int b200_mv_norm(const b200_mv* mv, int type, double* normvec, int nnorm) {
  if (!mv || !normvec) return fail(B200_ERR_ARG, "MvNorm: null argument");
  if (type != B200_NORM_ONE && type != B200_NORM_TWO && type != B200_NORM_INF)
    return fail(B200_ERR_ARG, "MvNorm: invalid norm type (valid: OneNorm, TwoNorm, InfNorm)");
  const int k = mv->ncols();
  if (nnorm < k) return 0;  // reference silently does nothing (BelosEpetraAdapter.cpp:516)
  if (k == 0) return 0;
  Ctx* ctx = mv->ctx;
  std::vector<uint64_t> ia(k);
  for (int j = 0; j < k; ++j) ia[j] = mv->id(j);
  if (type == B200_NORM_TWO && ctx->fuse) {  // x.x may already be known (fused epilogue / MvDot)
    bool known = true;
    for (int j = 0; j < k && known; ++j) {
      int rc = 0;
      double v = 0.0;
      known = ctx->recall_dot(ia[j], ia[j], &v, &rc);
      if (rc) return rc;
      if (known) normvec[j] = std::sqrt(v);
    }
    if (known) {
      ctx->dot_hits += k;
      ctx->answered_bytes += 8.0 * mv->n_local * k;
      return 0;
    }
  }
  B200_TRY(prep_read(mv));
  ++ctx->seq;
  ctx->plan_reduction(k, type == B200_NORM_INF);
  for (int j = 0; j < k;) {
    const int len = run_length(j, k, mv, j);
    B200_TRY(launch_norm(ctx, mv->col(j), mv->ld, mv->n_local, len, type, j, j + len >= k));
    j += len;
  }
  B200_TRY(ctx->finish_reduction(k, type == B200_NORM_INF));
  for (int j = 0; j < k; ++j) {
    if (type == B200_NORM_TWO) ctx->remember_dot(ia[j], ia[j], ctx->mbox_h[j]);
    normvec[j] = (type == B200_NORM_TWO) ? std::sqrt(ctx->mbox_h[j]) : ctx->mbox_h[j];
  }
  return 0;
}

extern "C" int b200_mv_trans_mv(const b200_mv* mv, double alpha, const b200_mv* A, double* B,
                                int ldb, int brows, int bcols) {
  if (!mv || !A || !B) return fail(B200_ERR_ARG, "MvTransMv: null argument");
  const int p = A->ncols(), q = mv->ncols();
  if (brows != p || bcols != q || ldb < std::max(1, brows))
    return failf(B200_ERR_ARG, "MvTransMv: B is %d x %d (ld %d) but A^T*mv is %d x %d", brows,
                 bcols, ldb, p, q);
  if (!same_map(mv, A)) return fail(B200_ERR_SHAPE, "MvTransMv: vector lengths differ");
  if (p == 0 || q == 0) return 0;
  if (alpha == 0.0) {  // DGEMM semantics: operands are not referenced
    for (int j = 0; j < q; ++j)
      for (int i = 0; i < p; ++i) B[(size_t)j * ldb + i] = 0.0;
    return 0;
  }
  // (the top of the mailbox, from kSpecBase, holds fused-epilogue results that pending memo entries point to)
  if ((size_t)p * q > kSpecBase)
    return fail(B200_ERR_ARG, "MvTransMv: result block larger than the host mailbox");
  Ctx* ctx = mv->ctx;
  // Is this the product the last (eligible) MvTimesMatAddMv could serve -- A^T V_new of the same columns?
  // The predictor learns from the answer whether or not that update speculated; if it did, the block waits
  // in the mailbox (update_gram.cu) and no kernel runs.
  if (ctx->fuse && ctx->spec.open && (int)ctx->spec.a_ids.size() == p && (int)ctx->spec.v_ids.size() == q &&
      !A->st->external && !mv->st->external) {
    bool same = true;
    for (int i = 0; i < p && same; ++i) same = ctx->spec.a_ids[(size_t)i] == A->st->ver[(size_t)A->cols[(size_t)i]];
    for (int j = 0; j < q && same; ++j) same = ctx->spec.v_ids[(size_t)j] == mv->st->ver[(size_t)mv->cols[(size_t)j]];
    if (same) {
      ctx->spec.open = false;
      ctx->spec.next_hit[ctx->spec.hist & 3u] = true;
      ctx->spec.hist = (ctx->spec.hist << 1) | 1u;
      Ctx::GramMemo& gm = ctx->gmemo;
      if (gm.valid && gm.a_ids == ctx->spec.a_ids && gm.v_ids == ctx->spec.v_ids) {
        B200_TRY(ctx->wait_flag(gm.seq));
        for (int j = 0; j < q; ++j)
          for (int i = 0; i < p; ++i) B[(size_t)j * ldb + i] = alpha * ctx->mbox_h[gm.src + (size_t)j * p + i];
        ++ctx->gram_hits;
        ctx->answered_bytes += 8.0 * mv->n_local * ((double)p + q);
        return 0;
      }
    }
  }
  // One vector against <= 8 columns (findBasis, BelosDGKSOrthoManager.hpp:1064): every entry may already be known --
  // X0 . Xj from the cross products of the block update, Xi . Xj (i >= 1) from the column step of Xi (findbasis.cu).
  if (ctx->fuse && q == 1 && p <= 8 && !A->st->external && !mv->st->external) {
    double vals[8];
    bool known = true;
    int rc = 0;
    for (int i = 0; i < p && known; ++i) known = ctx->recall_dot(A->id(i), mv->id(0), &vals[i], &rc);
    if (rc) return rc;
    if (known) {
      for (int i = 0; i < p; ++i) B[i] = alpha * vals[i];
      ++ctx->skinny_hits;
      ctx->answered_bytes += 8.0 * mv->n_local * ((double)p + 1.0);
      return 0;
    }
  }
  // otherwise the skinny kernel: it applies pending column scales itself and writes the scaled columns back -- the
  // normalisation of the previous column costs no launch
  double scales[9];
  bool fold = false;
  if (ctx->pending > 0 && gram_is_skinny(p, q) && run_length(0, p, A, 0) == p && mv->n_local > 0) {
    B200_TRY(prep_read(A, 0, p, true));
    B200_TRY(prep_read(mv, 0, 1, true));
    for (int i = 0; i < p; ++i) {
      scales[i] = A->st->pscale[(size_t)A->cols[(size_t)i]];
      fold = fold || scales[i] != 1.0;
    }
    scales[p] = mv->st->pscale[(size_t)mv->cols[0]];
    fold = fold || scales[p] != 1.0;
    // (the scaled columns are written back: copies of the stored, unscaled data cannot be pending -- MvScale
    // made them real before it deferred the factor)
  } else {
    B200_TRY(prep_read(A));
    B200_TRY(prep_read(mv));
  }
  ++ctx->seq;
  ctx->plan_reduction(p * q, false);
  if (mv->n_local == 0)  // empty row block: this rank contributes zeros (and still joins the exchange)
    B200_TRY(launch_zero_result(ctx, ctx->result_ptr(), p * q, ctx->make_pub(true)));
  for (int jv = 0; jv < q && mv->n_local > 0;) {
    const int lv = run_length(jv, q, mv, jv);
    for (int ja = 0; ja < p;) {
      const int la = run_length(ja, p, A, ja);
      const bool last = (jv + lv >= q) && (ja + la >= p);
      B200_TRY(launch_gram(ctx, A->col(ja), A->ld, la, mv->col(jv), mv->ld, lv, mv->n_local,
                           alpha, ja + jv * p, p, last, fold ? scales : nullptr));
      ja += la;
    }
    jv += lv;
  }
  if (fold) {  // the kernel stored the scaled columns
    for (int i = 0; i <= p; ++i) {
      Storage* st = i < p ? A->st.get() : mv->st.get();
      const int c = i < p ? A->cols[(size_t)i] : mv->cols[0];
      if (st->pscale[(size_t)c] != 1.0) {
        st->pscale[(size_t)c] = 1.0;
        --ctx->pending;
        ++ctx->scale_folded;
      }
    }
  }
  B200_TRY(ctx->finish_reduction(p * q, false));
  for (int j = 0; j < q; ++j)
    std::memcpy(B + (size_t)j * ldb, ctx->mbox_h + (size_t)j * p, sizeof(double) * p);
  return 0;
}

extern "C" int b200_mv_times_mat_add_mv(b200_mv* mv, double alpha, const b200_mv* A,
                                        const double* B, int ldb, int brows, int bcols,
                                        double beta) {
  if (!mv || !A || !B) return fail(B200_ERR_ARG, "MvTimesMatAddMv: null argument");
  const int p = A->ncols(), q = mv->ncols();
  if (brows != p || bcols != q || ldb < std::max(1, brows))
    return failf(B200_ERR_ARG, "MvTimesMatAddMv: B is %d x %d (ld %d) but A has %d and mv %d columns",
                 brows, bcols, ldb, p, q);
  if (!same_map(mv, A)) return fail(B200_ERR_SHAPE, "MvTimesMatAddMv: vector lengths differ");
  if (q == 0) return 0;
  Ctx* ctx = mv->ctx;
  if (alpha == 0.0 || p == 0) {
    if (beta == 0.0) return b200_mv_init(mv, 0.0);
    return b200_mv_scale(mv, beta);
  }
  if (columns_overlap(mv, A))
    return fail(B200_ERR_ALIAS, "MvTimesMatAddMv: A and mv share columns (DGEMM forbids aliasing)");
  // One column against <= 8 (a column step of findBasis): one kernel that also folds in the pending scales of A,
  // fills a pending lazy copy of the target and leaves the dot products of the steps to come (findbasis.cu).
  if (q == 1 && p <= 8 && ctx->fuse && fb_step_enabled() && run_length(0, p, A, 0) == p &&
      !A->st->external && !mv->st->external && (ctx->nranks == 1 || ctx->p2p)) {
    Storage* st = mv->st.get();
    const int c = mv->cols[0];
    B200_TRY(prep_read(A, 0, p, true));  // (pending copies INTO A's columns are made; their scales stay pending)
    B200_TRY(lazy_materialize(st, c));
    B200_TRY(scale_flush(st, c));
    double* save_old = nullptr;
    std::pair<Storage*, int> saved{nullptr, 0};
    if (beta != 0.0 && st->deps[(size_t)c].size() == 1) {
      saved = st->deps[(size_t)c][0];
      save_old = saved.first->colptr(saved.second);
    } else {
      B200_TRY(deps_materialize(st, c));
    }
    double coef[8], scales[8];
    for (int i = 0; i < p; ++i) {
      coef[i] = B[i];
      scales[i] = A->st->pscale[(size_t)A->cols[(size_t)i]];
    }
    // later columns of the block this column belongs to (the block the last multi-column update wrote)
    int nx = 0;
    if (ctx->last_block.uid == st->uid && c >= ctx->last_block.c0 && c < ctx->last_block.c0 + ctx->last_block.q) {
      nx = std::min(7, ctx->last_block.c0 + ctx->last_block.q - 1 - c);
      nx = std::max(0, std::min(nx, (int)st->ver.size() - 1 - c));
      for (int k = 1; k <= nx; ++k)
        if (st->lazy[(size_t)(c + k)].src || st->pscale[(size_t)(c + k)] != 1.0) nx = 0;  // (not plain data: skip)
    }
    // the step of column 1 (one previous column) also leaves X0 . Xk for the later columns, unless the block update
    // already did (block sizes 2 and 4: the cross products of update.cu)
    bool prev0_next = false;
    if (p == 1 && nx > 0) {
      double dummy;
      int rc = 0;
      prev0_next = !ctx->recall_dot(A->id(0), st->ver[(size_t)(c + 1)], &dummy, &rc);
      if (rc) return rc;
    }
    const int nent = 2 + nx + p + (prev0_next ? nx : 0);
    FusedDots fd;
    size_t src = 0;
    if (!fused_slot(ctx, &fd, &src, nent)) return fail(B200_ERR_ARG, "MvTimesMatAddMv: no fused slot (internal)");
    const uint64_t old_id = mv->id(0);
    if (mv->n_local > 0)
      B200_TRY(launch_fb_step(ctx, mv->col(0), A->col(0), A->ld, p, coef, scales, nx ? st->colptr(c + 1) : nullptr,
                              st->ld, nx, prev0_next, mv->n_local, alpha, beta, save_old, fd));
    else  // empty row block: this rank contributes zeros (and still joins the exchange)
      B200_TRY(launch_zero_result(ctx, fd.out, nent, fd.pub));
    mv->touch(0, 1);
    for (int i = 0; i < p; ++i) {  // the kernel stored the scaled columns
      Storage* sa = A->st.get();
      const int ca = A->cols[(size_t)i];
      if (sa->pscale[(size_t)ca] != 1.0) {
        sa->pscale[(size_t)ca] = 1.0;
        --ctx->pending;
        ++ctx->scale_folded;
      }
    }
    if (save_old) {  // the copy now holds the old column
      lazy_unlink(saved.first, saved.second);
      ++ctx->lazy_saved;
    }
    const uint64_t nid = mv->id(0);
    ctx->remember_dot(nid, nid, 0.0, fd.seq, src);
    if (beta != 0.0) ctx->remember_dot(nid, old_id, 0.0, fd.seq, src + 1);
    for (int k = 0; k < nx; ++k) ctx->remember_dot(nid, st->ver[(size_t)(c + 1 + k)], 0.0, fd.seq, src + 2 + (size_t)k);
    for (int i = 0; i < p; ++i) ctx->remember_dot(nid, A->id(i), 0.0, fd.seq, src + 2 + (size_t)nx + (size_t)i);
    if (prev0_next)
      for (int k = 0; k < nx; ++k)
        ctx->remember_dot(A->id(0), st->ver[(size_t)(c + 1 + k)], 0.0, fd.seq, src + 2 + (size_t)nx + (size_t)p + (size_t)k);
    return 0;
  }
  B200_TRY(prep_read(A));
  // One-column update whose target has exactly one pending lazy copy (findBasis: oldMXj): the kernel reads the old
  // values anyway and stores them into the copy on its way -- the CloneCopy launch never happens.
  double* save_old = nullptr;
  std::pair<Storage*, int> saved{nullptr, 0};
  if (ctx->pending > 0 && q == 1 && beta != 0.0 && mv->n_local > 0 && run_length(0, p, A, 0) == p) {
    Storage* st = mv->st.get();
    const int c = mv->cols[0];
    B200_TRY(lazy_materialize(st, c));
    B200_TRY(scale_flush(st, c));
    if (st->deps[(size_t)c].size() == 1) {
      saved = st->deps[(size_t)c][0];
      save_old = saved.first->colptr(saved.second);
    } else {
      B200_TRY(deps_materialize(st, c));
    }
  } else {
    B200_TRY(prep_write(mv, 0, q, beta == 0.0));
  }
  const size_t bytes = (size_t)p * q * sizeof(double);
  if (bytes > kStageSlotBytes)
    return fail(B200_ERR_ARG, "MvTimesMatAddMv: dense block larger than the staging slot");
  // B (compact p x q).  Small blocks that go to one of the q <= 4 kernels in ONE launch travel as a kernel parameter;
  // everything else is staged through the pinned ring and copied to the device in front of the launches.
  const bool b_param = run_length(0, p, A, 0) == p && run_length(0, q, mv, 0) == q && q <= 4 &&
                       (size_t)p * q <= (size_t)kBParamDoubles && mv->n_local > 0;
  const int slot = ctx->stage_next;
  ctx->stage_next = (slot + 1) % kStageSlots;
  if (ctx->stage_used[slot]) B200_CUDA(cudaEventSynchronize(ctx->stage_ev[slot]));
  double* hs = (double*)(ctx->stage_h + (size_t)slot * kStageSlotBytes);
  for (int j = 0; j < q; ++j)
    std::memcpy(hs + (size_t)j * p, B + (size_t)j * ldb, sizeof(double) * p);
  if (!b_param) {
    B200_CUDA(cudaMemcpyAsync(ctx->dB, hs, bytes, cudaMemcpyHostToDevice, ctx->stream));
    B200_CUDA(cudaEventRecord(ctx->stage_ev[slot], ctx->stream));
    ctx->stage_used[slot] = true;
  } else {
    ctx->stage_used[slot] = false;  // (the launch copies the parameter: the slot is free again at once)
  }
  // One launch covers the whole update (A and mv both plain strided blocks)?  Then the kernel
  // also emits x_new.x_new and x_new.x_old per column — the MvDot calls DGKS makes next.
  FusedDots fd;
  size_t src = 0;
  const bool plain = run_length(0, p, A, 0) == p && run_length(0, q, mv, 0) == q;
  // Fused update + Gram (update_gram.cu): when the call pattern says that MvTransMv(A, mv) comes next (the
  // second pass of block Gram-Schmidt, BelosDGKSOrthoManager.hpp:1436-1455), the kernel also leaves A^T mv_new
  // in the mailbox.  An eligible update that found no taker counts as a miss for the predictor.
  const bool eligible = plain && !save_old && !A->st->external && !mv->st->external &&
                        update_gram_fits(ctx, p, q, mv->n_global / std::max(1, ctx->nranks));  // (same answer on every rank)
  bool spec = false;
  if (eligible) {
    Ctx::SpecState& sp = ctx->spec;
    if (sp.open) {
      sp.next_hit[sp.hist & 3u] = false;
      sp.hist = sp.hist << 1;
    }
    spec = (sp.next_hit[sp.hist & 3u] || update_gram_always()) &&
           (size_t)p * q + 2 * (size_t)q <= (size_t)kSpecSlotDoubles &&
           fused_slot(ctx, &fd, &src, p * q + 2 * q, true);
  }
  // q <= 4 through one SIMT launch: also the cross products new_i . new_k, the `product` entries of the findBasis
  // column steps that follow (findbasis.cu)
  // (q = 2 and q = 4 are single launches of the SIMT kernel; q = 3 is split 2 + 1)
  const int ncross = (!spec && plain && (q == 2 || q == 4) && fb_step_enabled()) ? q * (q - 1) / 2 : 0;
  const bool fuse = spec || (plain && 2 * q + ncross <= kAuxSlotDoubles && fused_slot(ctx, &fd, &src, 2 * q + ncross));
  if (fuse && !spec) fd.cross = ncross;
  std::vector<uint64_t> old_id(q);
  for (int j = 0; j < q; ++j) old_id[j] = mv->id(j);
  if (spec && mv->n_local > 0)
    B200_TRY(launch_update_gram(ctx, mv->col(0), mv->ld, q, A->col(0), A->ld, p, mv->n_local, ctx->dB, p,
                                alpha, beta, fd, b_param ? hs : nullptr));
  for (int jv = 0; jv < q && !spec;) {
    const int lv = run_length(jv, q, mv, jv);
    bool first = true;
    for (int ja = 0; ja < p;) {
      const int la = run_length(ja, p, A, ja);
      B200_TRY(launch_update(ctx, mv->col(jv), mv->ld, lv, A->col(ja), A->ld, la, mv->n_local,
                             ctx->dB + ja + (size_t)jv * p, p, alpha, first ? beta : 1.0,
                             fuse ? &fd : nullptr, save_old, b_param ? hs : nullptr));
      first = false;
      ja += la;
    }
    jv += lv;
  }
  mv->touch(0, q);
  if (save_old) {  // the copy now holds the old column
    lazy_unlink(saved.first, saved.second);
    ++ctx->lazy_saved;
  }
  if (fuse && mv->n_local == 0)  // no kernel ran: the fused sums are zero, but must still be handed over
    B200_TRY(launch_zero_result(ctx, fd.out, spec ? p * q + 2 * q : 2 * q + ncross, fd.pub));
  if (fuse)
    for (int j = 0; j < q; ++j) {
      const uint64_t nid = mv->id(j);
      ctx->remember_dot(nid, nid, 0.0, fd.seq, src + 2 * j);
      if (beta != 0.0) ctx->remember_dot(nid, old_id[j], 0.0, fd.seq, src + 2 * j + 1);
    }
  if (fuse && !spec && ncross) {
    int e = 0;
    for (int i = 0; i < q; ++i)
      for (int k = i + 1; k < q; ++k, ++e) ctx->remember_dot(mv->id(i), mv->id(k), 0.0, fd.seq, src + 2 * (size_t)q + (size_t)e);
  }
  if (plain && q >= 2 && !mv->st->external) {
    ctx->last_block.uid = mv->st->uid;
    ctx->last_block.c0 = mv->cols[0];
    ctx->last_block.q = q;
  }
  if (eligible) {
    Ctx::SpecState& sp = ctx->spec;
    sp.a_ids.resize((size_t)p);
    sp.v_ids.resize((size_t)q);
    for (int i = 0; i < p; ++i) sp.a_ids[(size_t)i] = A->id(i);
    for (int j = 0; j < q; ++j) sp.v_ids[(size_t)j] = mv->id(j);
    sp.open = true;
    if (spec) {
      Ctx::GramMemo& gm = ctx->gmemo;
      gm.a_ids = sp.a_ids;
      gm.v_ids = sp.v_ids;
      gm.seq = fd.seq;
      gm.src = src + 2 * (size_t)q;
      gm.valid = true;
      ++ctx->gram_spec;
    }
  }
  return 0;
}

extern "C" int b200_mv_copy_to_host(const b200_mv* mv, double* host, int64_t ldh) {
  if (!mv || !host) return fail(B200_ERR_ARG, "copy_to_host: null argument");
  if (ldh < mv->n_local) return fail(B200_ERR_ARG, "copy_to_host: ldh < local length");
  Ctx* ctx = mv->ctx;
  B200_TRY(prep_read(mv));
  if (mv->n_local > 0) {
    for (int j = 0; j < mv->ncols();) {
      const int len = run_length(j, mv->ncols(), mv, j);
      B200_CUDA(cudaMemcpy2DAsync(host + (size_t)j * ldh, (size_t)ldh * 8, mv->col(j),
                                  (size_t)mv->ld * 8, (size_t)mv->n_local * 8, len,
                                  cudaMemcpyDeviceToHost, ctx->stream));
      j += len;
    }
  }
  B200_CUDA(cudaStreamSynchronize(ctx->stream));
  ++ctx->syncs;
  return 0;
}

extern "C" int b200_mv_copy_from_host(b200_mv* mv, const double* host, int64_t ldh) {
  if (!mv || !host) return fail(B200_ERR_ARG, "copy_from_host: null argument");
  if (ldh < mv->n_local) return fail(B200_ERR_ARG, "copy_from_host: ldh < local length");
  Ctx* ctx = mv->ctx;
  B200_TRY(prep_write(mv, 0, mv->ncols(), true));
  if (mv->n_local > 0) {
    for (int j = 0; j < mv->ncols();) {
      const int len = run_length(j, mv->ncols(), mv, j);
      B200_CUDA(cudaMemcpy2DAsync(mv->col(j), (size_t)mv->ld * 8, host + (size_t)j * ldh,
                                  (size_t)ldh * 8, (size_t)mv->n_local * 8, len,
                                  cudaMemcpyHostToDevice, ctx->stream));
      j += len;
    }
  }
  mv->touch(0, mv->ncols());
  B200_CUDA(cudaStreamSynchronize(ctx->stream));
  ++ctx->syncs;
  return 0;
}

extern "C" int b200_mv_device_ptr(const b200_mv* mv, int col, void** ptr, int64_t* ld) {
  if (!mv || col < 0 || col >= mv->ncols()) return fail(B200_ERR_ARG, "device_ptr: bad column");
  if (mv->ctx->pending > 0) {  // the caller sees raw memory from now on: nothing of this storage may stay pending
    Storage* st = mv->st.get();
    for (int c = 0; c < (int)st->ver.size(); ++c) {
      B200_TRY(deps_materialize(st, c));
      B200_TRY(lazy_materialize(st, c));
      B200_TRY(scale_flush(st, c));
    }
  }
  mv->st->external = true;  // the caller may write through the pointer: stop tracking contents
  if (ptr) *ptr = mv->col(col);
  if (ld) *ld = mv->ld;
  return 0;
}

// =========================================================================== operator
namespace {

template <class T>
int upload(T** dst, const std::vector<T>& src) {
  *dst = nullptr;
  // 512 B of zeroed slack: the SpMM bulk copies round segment ends up to 256-B multiples
  const size_t bytes = src.size() * sizeof(T) + 512;
  B200_CUDA(cudaMalloc((void**)dst, bytes));
  B200_CUDA(cudaMemset(*dst, 0, bytes));
  if (!src.empty())
    B200_CUDA(cudaMemcpy(*dst, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice));
  return 0;
}

}  // namespace

namespace {

void halo_window_release(b200_op* op) {
  for (size_t i = 0; i < op->peers.size() && i < (size_t)kMaxPeers; ++i)
    if (op->peer_hwin[i]) {
      cudaIpcCloseMemHandle(op->peer_hwin[i]);
      op->peer_hwin[i] = nullptr;
    }
  if (op->hwin) cudaFree(op->hwin);
  op->hwin = nullptr;
  if (op->p2p_halo) op->recv_buf = nullptr;  // it pointed into the window
  op->p2p_halo = false;
}

// (Re)allocate this rank's halo window for k right-hand sides and map the neighbours' windows.
// Collective: every rank reaches it in the same Apply (the block width is the same everywhere).
// Any failure on any rank, or a neighbour relation that is not mutual (a rank that only sends has
// no back-pressure on the parity scheme), leaves all ranks on the NCCL send/recv path.
int halo_window_setup(Ctx* ctx, b200_op* op, size_t k) {
  const int P = ctx->nranks, me = ctx->rank;
  B200_CUDA(cudaStreamSynchronize(ctx->stream));
  B200_CUDA(cudaStreamSynchronize(ctx->comm_stream));
  halo_window_release(op);
  int ok = ctx->p2p && op->peers.size() <= (size_t)kMaxPeers ? 1 : 0;
  for (const HaloPeer& pr : op->peers)
    if (pr.send_cnt <= 0 || pr.recv_cnt <= 0) ok = 0;
  const size_t half = ((size_t)op->ghost_ld * k * sizeof(double) + 255) / 256 * 256;
  const size_t bytes = 256 + 2 * half;
  cudaIpcMemHandle_t mine;
  std::memset(&mine, 0, sizeof(mine));
  if (ok && cudaMalloc(&op->hwin, bytes) != cudaSuccess) ok = 0;
  if (ok && cudaMemset(op->hwin, 0, bytes) != cudaSuccess) ok = 0;
  if (ok && cudaDeviceSynchronize() != cudaSuccess) ok = 0;  // zeroed before any peer can learn the handle
  if (ok && cudaIpcGetMemHandle(&mine, op->hwin) != cudaSuccess) ok = 0;
  cudaGetLastError();
  const size_t per = 9;
  std::vector<int64_t> h((size_t)P * per, 0);
  std::memcpy(&h[(size_t)me * per], &mine, 64);
  h[(size_t)me * per + 8] = ok;
  int64_t* d = nullptr;
  B200_CUDA(cudaMalloc((void**)&d, h.size() * sizeof(int64_t)));
  B200_CUDA(cudaMemcpy(d, h.data(), h.size() * sizeof(int64_t), cudaMemcpyHostToDevice));
  B200_TRY(nccl_allreduce_i64(ctx, d, h.size(), ctx->stream));
  B200_CUDA(cudaMemcpyAsync(h.data(), d, h.size() * sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
  B200_CUDA(cudaStreamSynchronize(ctx->stream));
  for (int r = 0; r < P; ++r) ok = ok && h[(size_t)r * per + 8] == 1;
  for (size_t i = 0; i < op->peers.size() && ok; ++i) {
    cudaIpcMemHandle_t hr;
    std::memcpy(&hr, &h[(size_t)op->peers[i].rank * per], 64);
    if (cudaIpcOpenMemHandle(&op->peer_hwin[i], hr, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
      cudaGetLastError();
      op->peer_hwin[i] = nullptr;
      ok = 0;
    }
  }
  int64_t v = ok;
  B200_CUDA(cudaMemcpy(d, &v, sizeof(v), cudaMemcpyHostToDevice));
  B200_TRY(nccl_allreduce_i64(ctx, d, 1, ctx->stream));
  B200_CUDA(cudaMemcpyAsync(&v, d, sizeof(v), cudaMemcpyDeviceToHost, ctx->stream));
  B200_CUDA(cudaStreamSynchronize(ctx->stream));
  cudaFree(d);
  if (v != P) {
    halo_window_release(op);
    return 0;
  }
  op->p2p_halo = true;
  op->hseq = 0;
  return 0;
}

}  // namespace

extern "C" int b200_op_create_csr(b200_ctx* ctx, int64_t n_global, int64_t row_begin,
                                  int64_t row_end, const int32_t* rowptr,
                                  const int64_t* colind_global, const double* values,
                                  b200_op** out) {
  if (!ctx || !rowptr || !out) return fail(B200_ERR_ARG, "b200_op_create_csr: null argument");
  int64_t rb, re;
  B200_TRY(b200_linear_partition(n_global, ctx->nranks, ctx->rank, &rb, &re));
  if (rb != row_begin || re != row_end)
    return failf(B200_ERR_ARG,
                 "b200_op_create_csr: rows [%lld,%lld) are not this rank's block [%lld,%lld) of "
                 "the linear map",
                 (long long)row_begin, (long long)row_end, (long long)rb, (long long)re);
  const int64_t n_local = re - rb;
  const int64_t nnz = rowptr[n_local];
  if (nnz > 0 && (!colind_global || !values))
    return fail(B200_ERR_ARG, "b200_op_create_csr: null column / value array");
  B200_CUDA(cudaSetDevice(ctx->device));

  std::vector<int32_t> col_local((size_t)nnz);
  std::vector<int64_t> ghost_gids((size_t)nnz);
  std::vector<int32_t> ghost_owner((size_t)nnz);
  const int64_t n_ghost = b200_build_colmap(n_global, ctx->nranks, ctx->rank, nnz, colind_global,
                                            col_local.data(), ghost_gids.data(),
                                            ghost_owner.data());
  if (n_ghost < 0) return B200_ERR_ARG;
  ghost_gids.resize((size_t)n_ghost);
  ghost_owner.resize((size_t)n_ghost);

  // rows sorted by local column id, as after Epetra_CrsGraph::SortIndices/OptimizeStorage
  std::vector<double> vals(values, values + nnz);
  std::vector<int32_t> rp(rowptr, rowptr + n_local + 1);
  std::vector<int32_t> bnd, perm;
  int max_row = 0;
  for (int64_t r = 0; r < n_local; ++r) {
    const int32_t b = rp[r], e = rp[r + 1];
    if (e < b) return fail(B200_ERR_ARG, "b200_op_create_csr: rowptr not monotone");
    max_row = std::max(max_row, e - b);
    bool sorted = true, ghost = false;
    for (int32_t k = b; k < e; ++k) {
      if (k > b && col_local[k] < col_local[k - 1]) sorted = false;
      if (col_local[k] >= n_local) ghost = true;
    }
    if (!sorted) {
      perm.resize(e - b);
      std::iota(perm.begin(), perm.end(), 0);
      std::stable_sort(perm.begin(), perm.end(),
                       [&](int a, int c) { return col_local[b + a] < col_local[b + c]; });
      std::vector<int32_t> tc(e - b);
      std::vector<double> tv(e - b);
      for (int i = 0; i < e - b; ++i) {
        tc[i] = col_local[b + perm[i]];
        tv[i] = vals[b + perm[i]];
      }
      std::copy(tc.begin(), tc.end(), col_local.begin() + b);
      std::copy(tv.begin(), tv.end(), vals.begin() + b);
    }
    if (ghost) bnd.push_back((int32_t)r);
  }

  auto* op = new b200_op();
  op->ctx = ctx;
  op->n_global = n_global;
  op->row_begin = rb;
  op->row_end = re;
  op->n_local = n_local;
  op->n_ghost = n_ghost;
  op->ghost_ld = ghost_pad(n_ghost);
  op->nnz = nnz;
  op->max_row_nnz = max_row;
  for (int i = 0; i < 3; ++i) {
    const int64_t R = 64 << i;
    int64_t m = 0;
    for (int64_t r0 = 0; r0 < n_local; r0 += R)
      m = std::max<int64_t>(m, rp[std::min(r0 + R, n_local)] - rp[r0]);
    op->max_tile_nnz[i] = m;
  }
  op->n_bnd = (int64_t)bnd.size();
  op->h_rowptr = rp;            // for the window plans of spmm_win.cu (built at the first apply)
  op->h_colind = col_local;
  op->h_values = vals;
  int rc = upload(&op->rowptr, rp);
  if (!rc) rc = upload(&op->colind, col_local);
  if (!rc) rc = upload(&op->values, vals);
  if (!rc) rc = upload(&op->bnd_rows, bnd);
  if (rc) {
    b200_op_destroy(op);
    return rc;
  }

  if (ctx->nranks > 1) {
    const int P = ctx->nranks, me = ctx->rank;
    // recv side: ghosts are sorted by GID, hence grouped by owner
    std::vector<int64_t> need(P, 0), recv_off(P, 0);
    rc = b200_halo_recv_plan(P, n_ghost, ghost_owner.data(), need.data(), recv_off.data());
    if (rc) {
      b200_op_destroy(op);
      return rc;
    }
    // everyone learns the full P x P "needs" matrix (row = requester)
    std::vector<int64_t> M((size_t)P * P, 0);
    for (int r = 0; r < P; ++r) M[(size_t)me * P + r] = need[r];
    int64_t* dM = nullptr;
    rc = upload(&dM, M);
    if (!rc) rc = nccl_allreduce_i64(ctx, dM, (size_t)P * P, ctx->stream);
    if (!rc && cudaMemcpyAsync(M.data(), dM, sizeof(int64_t) * P * P, cudaMemcpyDeviceToHost,
                               ctx->stream) != cudaSuccess)
      rc = fail(B200_ERR_CUDA, "halo setup: D2H copy failed");
    if (!rc && cudaStreamSynchronize(ctx->stream) != cudaSuccess)
      rc = fail(B200_ERR_CUDA, "halo setup: sync failed");
    cudaFree(dM);
    if (rc) {
      b200_op_destroy(op);
      return rc;
    }
    for (int64_t v : M) op->any_halo = op->any_halo || v > 0;  // M is identical on every rank
    std::vector<int64_t> send_cnt(P, 0), send_off(P, 0);
    for (int r = 0; r < P; ++r) send_cnt[r] = M[(size_t)r * P + me];
    for (int r = 1; r < P; ++r) send_off[r] = send_off[r - 1] + send_cnt[r - 1];
    const int64_t n_send = send_off[P - 1] + send_cnt[P - 1];
    // exchange the GID lists: I tell each owner which of its rows I need
    int64_t *d_need = nullptr, *d_asked = nullptr;
    rc = upload(&d_need, ghost_gids);
    std::vector<int64_t> asked((size_t)n_send, 0);
    if (!rc) rc = upload(&d_asked, asked);
    if (!rc) rc = nccl_group_start(ctx);
    for (int r = 0; r < P && !rc; ++r) {
      if (need[r] > 0)
        rc = nccl_send(ctx, d_need + recv_off[r], (size_t)need[r], false, r, ctx->stream);
      if (!rc && send_cnt[r] > 0)
        rc = nccl_recv(ctx, d_asked + send_off[r], (size_t)send_cnt[r], false, r, ctx->stream);
    }
    if (!rc) rc = nccl_group_end(ctx);
    if (!rc && n_send > 0 &&
        cudaMemcpyAsync(asked.data(), d_asked, sizeof(int64_t) * n_send, cudaMemcpyDeviceToHost,
                        ctx->stream) != cudaSuccess)
      rc = fail(B200_ERR_CUDA, "halo setup: D2H copy failed");
    if (!rc && cudaStreamSynchronize(ctx->stream) != cudaSuccess)
      rc = fail(B200_ERR_CUDA, "halo setup: sync failed");
    cudaFree(d_need);
    cudaFree(d_asked);
    if (rc) {
      b200_op_destroy(op);
      return rc;
    }
    std::vector<int32_t> send_idx((size_t)n_send);
    rc = b200_halo_send_index(n_global, P, me, n_send, asked.data(), send_idx.data());
    if (rc) {
      b200_op_destroy(op);
      return rc;
    }
    op->n_send = n_send;
    rc = upload(&op->send_idx, send_idx);
    if (rc) {
      b200_op_destroy(op);
      return rc;
    }
    for (int r = 0; r < P; ++r)
      if (need[r] > 0 || send_cnt[r] > 0) {
        HaloPeer hp{r, send_off[r], send_cnt[r], recv_off[r], need[r]};
        // rank r's ghosts are grouped by owner in ascending rank order (MakeColMap order)
        for (int q2 = 0; q2 < P; ++q2) {
          hp.remote_nghost += M[(size_t)r * P + q2];
          if (q2 < me) hp.remote_off += M[(size_t)r * P + q2];
        }
        op->peers.push_back(hp);
      }
  }
  // the uploads above went through the legacy default stream; the kernels run on non-blocking streams
  if (cudaDeviceSynchronize() != cudaSuccess) {
    b200_op_destroy(op);
    return fail(B200_ERR_CUDA, "b200_op_create_csr: device synchronisation failed");
  }
  *out = op;
  return 0;
}

extern "C" int b200_op_destroy(b200_op* op) {
  if (!op) return 0;
  cudaFree(op->rowptr);
  cudaFree(op->colind);
  cudaFree(op->values);
  cudaFree(op->bnd_rows);
  cudaFree(op->int_rows);
  cudaFree(op->send_idx);
  cudaFree(op->send_buf);
  halo_window_release(op);  // (p2p: recv_buf pointed into the window and is reset here)
  for (WinPlan& pl : op->win) free_win_plan(&pl);
  cudaFree(op->recv_buf);
  cudaFree(op->xcol);
  delete op;
  return 0;
}

extern "C" int b200_op_info(const b200_op* op, int64_t* n_local, int64_t* n_global,
                            int64_t* nnz_local, int64_t* n_ghost, int64_t* n_boundary_rows) {
  if (!op) return fail(B200_ERR_ARG, "b200_op_info: null operator");
  if (n_local) *n_local = op->n_local;
  if (n_global) *n_global = op->n_global;
  if (nnz_local) *nnz_local = op->nnz;
  if (n_ghost) *n_ghost = op->n_ghost;
  if (n_boundary_rows) *n_boundary_rows = op->n_bnd;
  return 0;
}

extern "C" int b200_op_apply(const b200_op* op_c, const b200_mv* x, b200_mv* y, int trans) {
  if (!op_c || !x || !y) return fail(B200_ERR_ARG, "Apply: null argument");
  if (trans != 0)
    return fail(B200_ERR_UNSUPPORTED, "Apply: transpose apply is not implemented on this path");
  b200_op* op = const_cast<b200_op*>(op_c);
  Ctx* ctx = op->ctx;
  if (x->ncols() != y->ncols())
    return fail(B200_ERR_ARG, "Apply: x and y have different numbers of vectors (Epetra: -2)");
  if (x->n_local != op->n_local || y->n_local != op->n_local || x->n_global != op->n_global)
    return fail(B200_ERR_SHAPE, "Apply: multivector map does not match the operator");
  const int k = x->ncols();
  if (k == 0) return 0;
  B200_TRY(prep_read(x));
  B200_TRY(prep_write(y, 0, k, !columns_overlap(x, y)));
  // X == Y: Epetra applies into a temporary (Epetra_CrsMatrix.cpp:3152); do the same
  b200_mv* xtmp = nullptr;
  if (columns_overlap(x, y)) {
    B200_TRY(b200_mv_clone_copy(x, nullptr, 0, &xtmp));
    x = xtmp;
  }
  int rc = 0;
  const double algo_bytes =
      12.0 * op->nnz + 4.0 * (op->n_local + 1) + 16.0 * (double)op->n_local * k;
  for (int j = 0; j < k && !rc;) {
    const int len = run_length(j, k, x, j, y, j);
    const double* xp = x->col(j);
    double* yp = y->col(j);
    if (ctx->nranks == 1 || !op->any_halo) {
      int wrc = 0;
      const WinPlan* pl = nullptr;
      const int kw = win_width(ctx, op, len, &pl, &wrc);
      if (wrc) { rc = wrc; break; }
      FusedDots fd;
      size_t src = 0;
      const bool fuse = len == k && k <= kAuxSlotDoubles && fused_slot(ctx, &fd, &src, k);
      ctx->t_begin(K_SPMM, algo_bytes * len / k);
      if (pl) {  // x staged in shared memory (spmm_win.cu), kw right-hand sides per launch
        const int g = spmm_win_grid(ctx, *pl, kw, pl->ntiles);
        for (int k0 = 0; k0 < len && !rc; k0 += kw) {
          FusedDots f2 = fd;
          f2.out = fd.out + k0;
          if (k0 + kw < len) f2.not_last();  // the last group hands the results over
          rc = launch_spmm_win(ctx, op, *pl, nullptr, pl->ntiles, xp + (size_t)k0 * x->ld, x->ld,
                               yp + (size_t)k0 * y->ld, y->ld, kw, k0, fuse ? &f2 : nullptr, g, 0, g, nullptr, nullptr);
        }
      } else if (op->n_local == 0 && fuse) {  // empty row block: the sums are zero, but must still be handed over
        rc = launch_zero_result(ctx, fd.out, k, fd.pub);
      } else {
        rc = launch_spmm(ctx, op, nullptr, 0, xp, x->ld, yp, y->ld, len, fuse ? &fd : nullptr);
      }
      ctx->t_end();
      y->touch(j, len);
      if (fuse && !rc)
        for (int c = 0; c < k; ++c) ctx->remember_dot(y->id(c), y->id(c), 0.0, fd.seq, src + c);
    } else {
      // grow halo buffers (collective: the block width is the same on every rank)
      if ((size_t)len > op->buf_k) {
        rc = halo_window_setup(ctx, op, (size_t)len);
        if (rc) break;
        if (op->p2p_halo) op->buf_k = (size_t)len;
      }
      if (op->p2p_halo) {
        // peer-memory halo: put || interior rows -> boundary rows (which wait for the flags
        // themselves); no NCCL call
        const uint64_t hs = ++op->hseq;
        const size_t half = ((size_t)op->ghost_ld * op->buf_k * sizeof(double) + 255) / 256 * 256;
        const size_t par = (size_t)(hs & 1u);
        HaloPut hp;
        HaloWait hw;
        hp.npeers = hw.n = (int)op->peers.size();
        hw.seq = hs;
        hw.err = ctx->flag_d + 1;
        for (int pi = 0; pi < hp.npeers; ++pi) {
          const HaloPeer& pr = op->peers[pi];
          const size_t rhalf = ((size_t)ghost_pad(pr.remote_nghost) * op->buf_k * sizeof(double) + 255) / 256 * 256;
          char* rw = static_cast<char*>(op->peer_hwin[pi]);
          hp.send_off[pi] = pr.send_off;
          hp.send_cnt[pi] = pr.send_cnt;
          hp.dst[pi] = reinterpret_cast<double*>(rw + 256 + par * rhalf) + pr.remote_off;
          hp.dst_ld[pi] = ghost_pad(pr.remote_nghost);
          hp.flag[pi] = reinterpret_cast<uint64_t*>(rw) + ctx->rank;
          hw.flag[pi] = reinterpret_cast<const uint64_t*>(op->hwin) + pr.rank;
        }
        op->recv_buf = reinterpret_cast<double*>(static_cast<char*>(op->hwin) + 256 + par * half);
        // the put runs on the comm stream next to the interior rows; nothing on the compute
        // stream waits for it except whatever follows this apply (x must stay intact until the
        // put has read it)
        if (cudaEventRecord(ctx->ev_compute, ctx->stream) != cudaSuccess ||
            cudaStreamWaitEvent(ctx->comm_stream, ctx->ev_compute, 0) != cudaSuccess) {
          rc = fail(B200_ERR_CUDA, "Apply: event handoff failed");
          break;
        }
        rc = launch_halo_put(ctx, hp, op->send_idx, op->n_send, xp, x->ld, len, hs);
        if (rc) break;
        cudaEventRecord(ctx->ev_comm, ctx->comm_stream);
        int wrc = 0;
        const WinPlan* pl = nullptr;
        const int kw = win_width(ctx, op, len, &pl, &wrc);
        if (wrc) { rc = wrc; break; }
        // every rank takes the fused y.y slot (the exchange behind it is collective), whichever
        // kernel its own rows use
        FusedDots fd;
        size_t src = 0;
        const bool fuse = len == k && k <= kAuxSlotDoubles && fused_slot(ctx, &fd, &src, k);
        ctx->t_begin(K_SPMM, algo_bytes * len / k);
        if (pl) {
          // ONE launch: the tiles without ghost runs first (next to the put on the comm stream), then the tiles
          // with ghost runs -- the kernel's producer warp waits for the neighbours' flags when it reaches them.
          // (Two launches -- interior, then boundary -- cost a kernel tail + ramp per apply: B200_SPMM_SPLIT_LAUNCH=1.)
          // (kw right-hand sides per launch; every group has its own partial-sum region)
          const int64_t n1 = pl->n_int_tiles, n2 = pl->n_bnd_tiles;
          static const bool split_launch = [] {
            const char* e = getenv("B200_SPMM_SPLIT_LAUNCH");
            return e && e[0] == '1';
          }();
          if (!split_launch) {
            const int g = spmm_win_grid(ctx, *pl, kw, n1 + n2);
            for (int k0 = 0; k0 < len && !rc; k0 += kw) {
              FusedDots f2 = fd;
              f2.out = fd.out + k0;
              f2.ws = fd.ws + (size_t)k0 * g;
              if (k0 + kw < len) f2.not_last();  // the last group exchanges the slot
              rc = launch_spmm_win(ctx, op, *pl, pl->mix_tiles_d, n1 + n2, xp + (size_t)k0 * x->ld, x->ld,
                                   yp + (size_t)k0 * y->ld, y->ld, kw, k0, fuse ? &f2 : nullptr, g, 0, g,
                                   n2 > 0 ? &hw : nullptr, nullptr, n1);
            }
          } else {
          const int g1 = n1 > 0 ? spmm_win_grid(ctx, *pl, kw, n1) : 0;
          const int g2 = n2 > 0 ? spmm_win_grid(ctx, *pl, kw, n2) : 0;
          for (int pass = 0; pass < 2 && !rc; ++pass)
            for (int k0 = 0; k0 < len && !rc; k0 += kw) {
              FusedDots f2 = fd;
              f2.out = fd.out + k0;
              f2.ws = fd.ws + (size_t)k0 * (g1 + g2);
              if (!(k0 + kw >= len && (pass == 1 || n2 == 0))) f2.not_last();  // the very last launch exchanges the slot
              const double* xk = xp + (size_t)k0 * x->ld;
              double* yk = yp + (size_t)k0 * y->ld;
              if (pass == 0 && n1 > 0)
                rc = launch_spmm_win(ctx, op, *pl, pl->int_tiles_d, n1, xk, x->ld, yk, y->ld, kw, k0,
                                     fuse ? &f2 : nullptr, g1 + g2, 0, n2 > 0 ? 0 : g1, nullptr, nullptr);
              if (pass == 1 && n2 > 0)
                rc = launch_spmm_win(ctx, op, *pl, pl->bnd_tiles_d, n2, xk, x->ld, yk, y->ld, kw, k0,
                                     fuse ? &f2 : nullptr, g1 + g2, g1, g1 + g2, &hw, nullptr);
            }
          }
        } else {
          rc = launch_spmm(ctx, op, nullptr, -1, xp, x->ld, yp, y->ld, len);
          if (!rc) rc = launch_spmm_boundary(ctx, op, hw, xp, x->ld, yp, y->ld, len);
          if (!rc && fuse) {
            // gather kernels: no epilogue in the split apply -- one reduction launch into the fused slot, then
            // the exchange as a kernel of its own
            rc = launch_dot_into(ctx, yp, y->ld, op->n_local, len, fd.out);
            if (!rc) rc = fused_finish_separately(ctx, fd);
          }
        }
        ctx->t_end();
        if (rc) break;
        cudaStreamWaitEvent(ctx->stream, ctx->ev_comm, 0);
        y->touch(j, len);
        if (fuse && !rc)
          for (int c = 0; c < k; ++c) ctx->remember_dot(y->id(c), y->id(c), 0.0, fd.seq, src + c);
        j += len;
        continue;
      }
      if ((size_t)len > op->buf_k) {
        cudaFree(op->send_buf);
        cudaFree(op->recv_buf);
        op->send_buf = op->recv_buf = nullptr;
        if (cudaMalloc((void**)&op->send_buf,
                       std::max<size_t>(1, (size_t)op->n_send * len) * 8) != cudaSuccess ||
            cudaMalloc((void**)&op->recv_buf,
                       std::max<size_t>(1, (size_t)op->ghost_ld * len) * 8) != cudaSuccess) {
          rc = fail(B200_ERR_CUDA, "Apply: halo buffer allocation failed");
          break;
        }
        op->buf_k = (size_t)len;
      }
      // 1. pack the rows my neighbours need (compute stream: x is produced there)
      ctx->t_begin(K_HALO, 16.0 * (double)op->n_send * len);
      rc = launch_pack(ctx, ctx->stream, op->send_idx, op->n_send, xp, x->ld, len, op->send_buf);
      ctx->t_end();
      if (rc) break;
      if (cudaEventRecord(ctx->ev_compute, ctx->stream) != cudaSuccess ||
          cudaStreamWaitEvent(ctx->comm_stream, ctx->ev_compute, 0) != cudaSuccess) {
        rc = fail(B200_ERR_CUDA, "Apply: event handoff failed");
        break;
      }
      // 2. NCCL send/recv on the comm stream, one slice per (peer, rhs) so the ghost block is
      //    received directly in its final [rhs][ghost] layout (no unpack kernel)
      rc = nccl_group_start(ctx);
      for (size_t pi = 0; pi < op->peers.size() && !rc; ++pi) {
        const HaloPeer& pr = op->peers[pi];
        for (int c = 0; c < len && !rc; ++c) {
          if (pr.send_cnt > 0)
            rc = nccl_send(ctx, op->send_buf + (size_t)c * op->n_send + pr.send_off,
                           (size_t)pr.send_cnt, true, pr.rank, ctx->comm_stream);
          if (!rc && pr.recv_cnt > 0)
            rc = nccl_recv(ctx, op->recv_buf + (size_t)c * op->ghost_ld + pr.recv_off,
                           (size_t)pr.recv_cnt, true, pr.rank, ctx->comm_stream);
        }
      }
      if (!rc) rc = nccl_group_end(ctx);
      if (rc) break;
      ++ctx->launches;
      cudaEventRecord(ctx->ev_comm, ctx->comm_stream);
      // 3. interior rows while the halo is in flight
      ctx->t_begin(K_SPMM, algo_bytes * len / k);
      rc = launch_spmm(ctx, op, nullptr, -1, xp, x->ld, yp, y->ld, len);
      if (rc) break;
      // 4. boundary rows once the ghosts have landed
      cudaStreamWaitEvent(ctx->stream, ctx->ev_comm, 0);
      rc = launch_spmm(ctx, op, op->bnd_rows, op->n_bnd, xp, x->ld, yp, y->ld, len);
      ctx->t_end();
      y->touch(j, len);
    }
    j += len;
  }
  if (xtmp) b200_mv_destroy(xtmp);
  return rc;
}

// =========================================================================== preconditioning
extern "C" int b200_mv_multiply_elementwise(b200_mv* mv, double scalar_ab, const b200_mv* A,
                                            const b200_mv* B, double scalar_this) {
  if (!mv || !A || !B) return fail(B200_ERR_ARG, "Multiply: null multivector");
  if (A->ncols() != 1)
    return fail(B200_ERR_ARG, "Multiply: A must be a single vector (broadcast over the columns)");
  if (B->ncols() != mv->ncols())
    return fail(B200_ERR_ARG, "Multiply: number of vectors differs (Epetra Multiply: -1)");
  if (!same_map(mv, A) || !same_map(mv, B))
    return fail(B200_ERR_SHAPE, "Multiply: vector lengths differ (Epetra Multiply: -2)");
  Ctx* ctx = mv->ctx;
  const int k = mv->ncols();
  B200_TRY(prep_read(A));
  B200_TRY(prep_read(B));
  B200_TRY(prep_write(mv, 0, k, false));
  for (int j = 0; j < k;) {
    const int len = run_length(j, k, mv, j, B, j);
    B200_TRY(launch_ewise_mul(ctx, mv->col(j), mv->ld, A->col(0), B->col(j), B->ld, mv->n_local, len,
                              scalar_ab, scalar_this));
    mv->touch(j, len);
    j += len;
  }
  return 0;
}

extern "C" int b200_op_extract_diagonal(const b200_op* op, b200_mv** diag) {
  if (!op || !diag) return fail(B200_ERR_ARG, "ExtractDiagonalCopy: null argument");
  B200_TRY(new_mv(op->ctx, op->n_local, op->n_global, op->row_begin, 1, diag));
  return launch_diag(op->ctx, op, (*diag)->col(0), false, 0.0);
}

struct b200_cheby {
  b200_op* op = nullptr;
  b200_cheby_params p;
  b200_mv* invd = nullptr;      // guarded reciprocal of the diagonal
  double lambda_max = 0.0, lambda_min = 0.0;
  b200_mv *W = nullptr, *V = nullptr;  // work blocks, kept for the last block width
};

extern "C" void b200_cheby_default_params(b200_cheby_params* p) {
  if (!p) return;
  p->degree = 1;
  p->lambda_max = -1.0;
  p->eig_ratio = 30.0;
  p->min_diag = 0.0;
  p->eig_iters = 10;
  p->seed = 1;
}

extern "C" int b200_cheby_destroy(b200_cheby* c) {
  if (!c) return 0;
  b200_mv_destroy(c->invd);
  b200_mv_destroy(c->W);
  b200_mv_destroy(c->V);
  delete c;
  return 0;
}

extern "C" int b200_cheby_create(b200_op* op, const b200_cheby_params* p, b200_cheby** out) {
  if (!op || !p || !out) return fail(B200_ERR_ARG, "b200_cheby_create: null argument");
  if (p->degree <= 0)  // Ifpack_Chebyshev::Compute: "at least one application" (-2)
    return fail(B200_ERR_ARG, "b200_cheby_create: polynomial degree must be >= 1");
  Ctx* ctx = op->ctx;
  auto* c = new b200_cheby();
  c->op = op;
  c->p = *p;
  int rc = new_mv(ctx, op->n_local, op->n_global, op->row_begin, 1, &c->invd);
  if (!rc) rc = launch_diag(ctx, op, c->invd->col(0), true, p->min_diag);
  double lam = p->lambda_max;
  if (!rc && lam < 0) {
    // Ifpack_Chebyshev::PowerMethod (:729-759) on D^-1 A, start vector from the Epetra LCG
    b200_mv *x = nullptr, *y = nullptr;
    rc = new_mv(ctx, op->n_local, op->n_global, op->row_begin, 1, &x);
    if (!rc) rc = new_mv(ctx, op->n_local, op->n_global, op->row_begin, 1, &y);
    const unsigned int saved = ctx->seed;
    if (!rc) rc = b200_ctx_set_seed(static_cast<b200_ctx*>(ctx), p->seed);
    if (!rc) rc = b200_mv_random(x);
    ctx->seed = saved;
    double norm = 0.0, top = 0.0, bot = 0.0;
    if (!rc) rc = b200_mv_norm(x, B200_NORM_TWO, &norm, 1);
    if (!rc && norm == 0.0) rc = fail(B200_ERR_ARG, "power method: zero start vector");
    if (!rc) rc = b200_mv_scale(x, 1.0 / norm);
    lam = 0.0;
    for (int it = 0; it < p->eig_iters && !rc; ++it) {
      rc = b200_op_apply(op, x, y, 0);
      if (!rc) rc = b200_mv_multiply_elementwise(y, 1.0, c->invd, y, 0.0);
      if (!rc) rc = b200_mv_dot(y, x, &top, 1);
      if (!rc) rc = b200_mv_dot(x, x, &bot, 1);
      lam = top / bot;
      if (!rc) rc = b200_mv_norm(y, B200_NORM_TWO, &norm, 1);
      if (!rc && norm == 0.0) rc = fail(B200_ERR_ARG, "power method: zero iterate");
      if (!rc) rc = b200_mv_add_mv(x, 1.0 / norm, y, 0.0, y);
    }
    b200_mv_destroy(x);
    b200_mv_destroy(y);
  }
  if (rc) {
    b200_cheby_destroy(c);
    return rc;
  }
  // Compute(): "Test for Exact Preconditioned case"
  if (std::fabs(lam - 1.0) < 1e-6) {
    c->lambda_max = c->lambda_min = 1.0;
  } else {
    c->lambda_max = lam;
    c->lambda_min = lam / p->eig_ratio;
  }
  *out = c;
  return 0;
}

extern "C" int b200_cheby_lambda(const b200_cheby* c, double* lambda_max, double* lambda_min) {
  if (!c) return fail(B200_ERR_ARG, "b200_cheby_lambda: null smoother");
  if (lambda_max) *lambda_max = c->lambda_max;
  if (lambda_min) *lambda_min = c->lambda_min;
  return 0;
}

extern "C" int b200_cheby_apply(b200_cheby* c, const b200_mv* x, b200_mv* y) {
  if (!c || !x || !y) return fail(B200_ERR_ARG, "b200_cheby_apply: null argument");
  if (x->ncols() != y->ncols()) return fail(B200_ERR_ARG, "ApplyInverse: number of vectors differs (-2)");
  if (!same_map(x, y) || x->n_local != c->op->n_local)
    return fail(B200_ERR_SHAPE, "ApplyInverse: vector lengths differ from the operator's map");
  Ctx* ctx = x->ctx;
  const int k = x->ncols();
  if (k == 0) return 0;
  B200_TRY(prep_read(x));
  B200_TRY(prep_write(y, 0, k, !columns_overlap(x, y)));
  // AztecOO-style aliasing (x and y the same memory): work on a copy of x, as :482-488
  b200_mv* xcopy = nullptr;
  if (columns_overlap(x, y)) {
    B200_TRY(b200_mv_clone_copy(x, nullptr, 0, &xcopy));
    x = xcopy;
  }
  int rc = 0;
  const double* invd = c->invd->col(0);
  const int64_t n = x->n_local;
  if (c->lambda_min == 1.0 && c->lambda_max == c->lambda_min) {  // identity case :505-525
    rc = b200_mv_multiply_elementwise(y, 1.0, c->invd, x, 0.0);
    if (xcopy) b200_mv_destroy(xcopy);
    return rc;
  }
  if (!c->W || c->W->ncols() != k) {
    b200_mv_destroy(c->W);
    b200_mv_destroy(c->V);
    c->W = c->V = nullptr;
    rc = new_mv(ctx, x->n_local, x->n_global, x->row_offset, k, &c->W);
    if (!rc) rc = new_mv(ctx, x->n_local, x->n_global, x->row_offset, k, &c->V);
  }
  // coefficients :529-533 (delta is the inverse of ML_Cheby's)
  const double alpha = c->lambda_max / c->p.eig_ratio;
  const double beta = 1.1 * c->lambda_max;
  const double delta = 2.0 / (beta - alpha);
  const double theta = 0.5 * (beta + alpha);
  const double s1 = theta * delta;
  const double oot = 1.0 / theta;
  for (int j = 0; j < k && !rc;) {
    const int len = run_length(j, k, x, j, y, j);
    rc = launch_cheby_first(ctx, c->W->col(j), c->W->ld, y->col(j), y->ld, invd, x->col(j), x->ld, n,
                            len, oot, k == 1);
    j += len;
  }
  y->touch(0, k);
  double rhok = 1.0 / s1;
  for (int it = 0; it < c->p.degree - 1 && !rc; ++it) {
    rc = b200_op_apply(c->op, y, c->V, 0);
    const double rhokp1 = 1.0 / (2.0 * s1 - rhok);
    const double d1 = rhokp1 * rhok, d2 = 2.0 * rhokp1 * delta;
    rhok = rhokp1;
    for (int j = 0; j < k && !rc;) {
      const int len = run_length(j, k, x, j, y, j);
      rc = launch_cheby_step(ctx, c->W->col(j), c->W->ld, y->col(j), y->ld, invd, x->col(j), x->ld,
                             c->V->col(j), c->V->ld, n, len, d1, d2);
      j += len;
    }
    y->touch(0, k);
  }
  if (xcopy) b200_mv_destroy(xcopy);
  return rc;
}
