/*
 * b200_belos.h — C-ABI of the B200-native Block GMRES hot path.
 *
 * This is the drop-in boundary underneath Belos' plugin surface: the C++ shim
 * include/B200BelosAdapter.hpp (B200MultiVec : Belos::MultiVec<double>,
 * B200CrsOp : Belos::Operator<double>) forwards every virtual to one of the
 * entry points below.  Each entry point cites the reference interface it
 * replaces (paths relative to the reference tree, packages/...).
 *
 * Conventions (mirroring Epetra, epetra/src/Epetra_ConfigDefs.h EPETRA_CHK_ERR):
 *   - every function returns int: 0 = ok, <0 = error, >0 = warning;
 *     b200_last_error() returns a static, thread-local message for the last
 *     non-zero return.  The C++ shim turns non-zero into a Belos exception.
 *   - handles are opaque; the caller owns them and must destroy them.
 *   - all multivector storage is FP64, column-major, one leading dimension
 *     per allocation (ld >= n_local, padded to 32 elements).
 *   - every call is synchronous from the caller's point of view when it
 *     returns host data (dot / norm / trans_mv / copy_to_host); all other
 *     calls are stream-ordered on the context's compute stream.
 *   - no CPU fallback: without a CUDA device every compute entry point
 *     returns B200_ERR_NO_DEVICE.
 */
#ifndef B200_BELOS_H
#define B200_BELOS_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_OK 0
#define B200_ERR_ARG (-1)        /* bad argument (Epetra: -1 / -2 size mismatch) */
#define B200_ERR_SHAPE (-2)
#define B200_ERR_CUDA (-3)
#define B200_ERR_NO_DEVICE (-4)
#define B200_ERR_NCCL (-5)
#define B200_ERR_ALIAS (-6)
#define B200_ERR_UNSUPPORTED (-7)
#define B200_ERR_TIMEOUT (-8)    /* a reduction result or a peer GPU's exchange flag did not arrive (120 s / 30 s) */
#define B200_ERR_BELOS (-9)      /* a Belos exception that did not come from a C-ABI status (orthogonalisation
                                    failure, LinearProblem / status-test errors); solver library only */

/* Belos::NormType (belos/src/BelosTypes.hpp) */
#define B200_NORM_ONE 0
#define B200_NORM_TWO 1
#define B200_NORM_INF 2

typedef struct b200_ctx b200_ctx; /* device, streams, pools, mailboxes, NCCL comm */
typedef struct b200_mv b200_mv;   /* multivector (owning, or view sharing storage) */
typedef struct b200_op b200_op;   /* row-block distributed CSR operator */

const char* b200_last_error(void);
/* status code that came with b200_last_error() (thread-local; 0 if no call has failed yet) */
int b200_last_status(void);
const char* b200_version(void);

/* ------------------------------------------------------------------ context
 * Replaces Epetra_Comm (epetra/src/Epetra_SerialComm.cpp:117-152,
 * epetra/src/Epetra_MpiComm.cpp:62,126,154): one context per process/GPU. */
int b200_ctx_create(int device, b200_ctx** out);
int b200_ctx_destroy(b200_ctx* ctx);
int b200_ctx_sync(b200_ctx* ctx);
/* NCCL communicator (replaces MPI_Comm in Epetra_MpiComm). unique id = 128 bytes. */
int b200_comm_unique_id(void* id128);
int b200_ctx_comm_init(b200_ctx* ctx, const void* id128, int nranks, int rank);
int b200_ctx_rank(const b200_ctx* ctx, int* rank, int* nranks);
/* seed of the Park-Miller generator behind MvRandom (epetra/src/Epetra_Util.cpp:71-99) */
int b200_ctx_set_seed(b200_ctx* ctx, unsigned int seed);
/* counters: kernels launched / host syncs since creation (bench "gpu_launches") */
int b200_ctx_counters(const b200_ctx* ctx, int64_t* launches, int64_t* syncs);
/* Fused reductions: kernels that write a column (operator apply, MvTimesMatAddMv) also emit
 * the column inner products Belos' orthogonalisation asks for next, and MvDot / MvNorm(TwoNorm) on columns
 * whose contents have not changed since their product was computed return it without a launch.  Contents
 * are tracked by per-column ids; b200_mv_device_ptr() switches tracking off for that storage.
 * on = 0 disables both (every MvDot / MvNorm launches its own reduction); also B200_NO_FUSE=1 in the
 * environment.  dot_hits (may be NULL) returns how many column products were served that way. */
int b200_ctx_set_fusion(b200_ctx* ctx, int on);
int b200_ctx_fusion_stats(const b200_ctx* ctx, int64_t* dot_hits);
/* Fused update + Gram: block Gram-Schmidt makes two passes, C = Q^T X; X -= Q C; C2 = Q^T X; X -= Q C2
 * (belos/src/BelosDGKSOrthoManager.hpp:1400-1472).  When the call pattern predicts that MvTransMv(Q, X) follows an
 * MvTimesMatAddMv(-1, Q, C, 1, X), the update kernel also accumulates Q^T X_new from the tile of Q it has staged and the
 * MvTransMv is answered from the mailbox: one pass over Q of four disappears.  launches = updates that ran fused,
 * hits = MvTransMv calls answered that way.  B200_SPEC_GRAM=0 in the environment disables it, =2 fuses every
 * eligible update (p <= 256 basis columns, <= 8 right-hand sides). */
int b200_ctx_spec_gram_stats(const b200_ctx* ctx, int64_t* launches, int64_t* hits);
/* findBasis column steps: the one-column MvTimesMatAddMv of DGKS findBasis also emits the products of
 * the new column with the later columns of its block, and the block update before it the cross products of its columns,
 * so that the one-column MvTransMv calls of the following column steps (belos/src/BelosDGKSOrthoManager.hpp:1064) are
 * answered without a launch.  gram_calls_answered = such calls.  B200_FB_STEP=0 in the environment disables it. */
int b200_ctx_findbasis_stats(const b200_ctx* ctx, int64_t* gram_calls_answered);
/* SURVEY 8(d) algorithmic bytes of the MvTransMv / MvDot / MvNorm calls that were answered from those memos, i.e.
 * without a kernel of their own, since creation or the last reset (bench: whole-solve byte model). */
int b200_ctx_answered_bytes(b200_ctx* ctx, double* bytes, int reset);
/* Launches that never happened: single-column CloneCopy calls whose copy was made by the update kernel on its way
 * or was never needed (lazy copies), and single-column MvScale factors applied by the skinny Gram kernel that read
 * the column next (deferred scales).  DGKS findBasis makes one of each per column
 * (belos/src/BelosDGKSOrthoManager.hpp:1037, :1217-1222). */
int b200_ctx_deferral_stats(const b200_ctx* ctx, int64_t* lazy_copies_saved, int64_t* scales_folded);
/* per-kernel-class device timing (CUDA events on the compute stream).
 * classes: 0 spmm, 1 trans_mv, 2 times_mat_add_mv, 3 dot/norm, 4 axpby/scale/copy, 5 halo,
 *          6 times_mat_add_mv fused with the next trans_mv (update_gram.cu; bytes = those of the update alone),
 *          7 findBasis column step (findbasis.cu) */
#define B200_NUM_KCLASS 8
int b200_ctx_timing_enable(b200_ctx* ctx, int on);
int b200_ctx_timing_read(b200_ctx* ctx, double* ms, int64_t* calls, double* bytes, int reset);
/* raw stream handle (cudaStream_t) so a host harness can record its own events */
void* b200_ctx_stream(b200_ctx* ctx);
/* write >L2-sized scratch buffer (bench hygiene between timed iterations) */
int b200_ctx_flush_l2(b200_ctx* ctx);

/* -------------------------------------------------------------- multivector
 * Replaces Belos::EpetraMultiVec / Epetra_MultiVector
 * (belos/epetra/src/BelosEpetraAdapter.cpp:318-570,
 *  epetra/src/Epetra_MultiVector.cpp). */
/* ctor: Epetra_MultiVector(Map, numvecs, zeroOut)  Epetra_MultiVector.cpp:60-80 */
int b200_mv_create(b200_ctx* ctx, int64_t n_local, int64_t n_global, int64_t row_offset,
                   int ncols, int zero_out, b200_mv** out);
int b200_mv_destroy(b200_mv* mv);
/* Clone: same map, numvecs columns, UNINITIALISED   BelosEpetraAdapter.cpp:361-365 */
int b200_mv_clone(const b200_mv* mv, int ncols, b200_mv** out);
/* CloneCopy(index) (index==NULL: all columns)       BelosEpetraAdapter.cpp:368-379 */
int b200_mv_clone_copy(const b200_mv* mv, const int* index, int nindex, b200_mv** out);
/* CloneView / CloneViewNonConst(index): shares storage; index may be unordered
 * and repeated (belos/src/BelosMVOPTester.hpp:197-201)  BelosEpetraAdapter.cpp:382-393 */
int b200_mv_clone_view(const b200_mv* mv, const int* index, int nindex, b200_mv** out);
int64_t b200_mv_global_length(const b200_mv* mv); /* GetGlobalLength */
int64_t b200_mv_local_length(const b200_mv* mv);
int b200_mv_num_vecs(const b200_mv* mv);          /* GetNumberVecs   */
b200_ctx* b200_mv_ctx(const b200_mv* mv);
/* mv <- alpha*A*B + beta*mv ; B is a host column-major brows x bcols block with
 * leading dimension ldb (Teuchos::SerialDenseMatrix::values()/stride()).
 * BelosEpetraAdapter.cpp:421-435 -> Epetra_MultiVector::Multiply('N','N') */
int b200_mv_times_mat_add_mv(b200_mv* mv, double alpha, const b200_mv* A, const double* B,
                             int ldb, int brows, int bcols, double beta);
/* mv <- alpha*A + beta*B (A, B may alias mv)   BelosEpetraAdapter.cpp:437-452
 * -> Epetra_MultiVector::Update(a,A,b,B,0.0)  Epetra_MultiVector.cpp:1341-1500 */
int b200_mv_add_mv(b200_mv* mv, double alpha, const b200_mv* A, double beta, const b200_mv* B);
/* MvScale(alpha), MvScale(vector)              BelosEpetraAdapter.cpp:454-478 */
int b200_mv_scale(b200_mv* mv, double alpha);
int b200_mv_scale_cols(b200_mv* mv, const double* alpha, int nalpha);
/* B(brows x bcols, ldb) <- alpha * A^T * mv ; honours ldb, never resizes.
 * BelosEpetraAdapter.cpp:480-495 -> Multiply('T','N') + Reduce()
 * (Epetra_MultiVector.cpp:1936-2083,2400-2433).  Multi-rank: NCCL allreduce. */
int b200_mv_trans_mv(const b200_mv* mv, double alpha, const b200_mv* A, double* B, int ldb,
                     int brows, int bcols);
/* b[j] <- sum_i mv(i,j)*A(i,j)                 BelosEpetraAdapter.cpp:497-513 */
int b200_mv_dot(const b200_mv* mv, const b200_mv* A, double* b, int nb);
/* norms (type = B200_NORM_*)                    BelosEpetraAdapter.cpp:515-545,
 * Epetra_MultiVector.cpp:1508-1545 (Norm1), 1547-1577 (Norm2), 1580-1630 (NormInf) */
int b200_mv_norm(const b200_mv* mv, int type, double* normvec, int nnorm);
/* SetBlock(A, index): mv[:, index[i]] <- A[:, i]  BelosEpetraAdapter.cpp:395-419 */
int b200_mv_set_block(b200_mv* mv, const b200_mv* A, const int* index, int nindex);
/* MvRandom: Park-Miller LCG in (-1,1), sequence index = global_row + col*n_global
 * Epetra_MultiVector.cpp:490-515, Epetra_Util.cpp:71-99 */
int b200_mv_random(b200_mv* mv);
/* MvInit(alpha) = PutScalar                      Epetra_MultiVector.cpp:595-608 */
int b200_mv_init(b200_mv* mv, double alpha);
/* host <-> device transfer of the local rows (ExtractCopy / ctor from array) */
/* TSQR hooks of Belos::MultiVec (belos/src/BelosMultiVec.hpp:213-293), SURVEY 8 f.3 -- what
 * Belos::TsqrOrthoManagerImpl calls through MultiVecTraits<>::tsqr_adaptor_type
 * (belos/src/BelosTsqrOrthoManagerImpl.hpp:1546-1550).
 * factor_explicit: A = Q R with orthonormal Q (n x b, another multivector of the same map) and the b x b factor R
 *   (column-major, leading dimension ldr; upper triangular with a positive diagonal unless A is numerically rank
 *   deficient, in which case R = Q^T A is full and Q's trailing columns span a complement -- reveal_rank sorts that
 *   out).  Cholesky QR on the Gram / update kernels (trilinos_b200/csrc/tsqr.cu); b <= 64.
 * reveal_rank: SVD of R = U S V^T (host); *rank = number of singular values > tol * largest.  Full rank: Q and R
 *   unchanged; otherwise Q := Q U, R := S V^T (BelosMultiVec.hpp:252-283). */
int b200_mv_factor_explicit(const b200_mv* A, b200_mv* Q, double* R, int ldr, int force_nonnegative_diagonal);
int b200_mv_reveal_rank(b200_mv* Q, double* R, int ldr, double tol, int* rank);

int b200_mv_copy_to_host(const b200_mv* mv, double* host, int64_t ldh);
int b200_mv_copy_from_host(b200_mv* mv, const double* host, int64_t ldh);
/* raw device pointer of column j and the leading dimension (harness / tests) */
int b200_mv_device_ptr(const b200_mv* mv, int col, void** ptr, int64_t* ld);

/* out[j] <- sum_i x(i,j)*y(i,j), j < ncols: the local arithmetic of b200_mv_dot WITHOUT any allocation.
 * No counterpart in the reference: added for callers that cannot
 * call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Needs no context.
 * x, y (column-major, leading dimensions ldx, ldy >= n) and out (ncols doubles) are DEVICE pointers on the
 * caller's current device; stream is a cudaStream_t (NULL = default stream); the call is stream-ordered and
 * does not synchronise.  The per-CTA partial sums live in a static __device__ array of 262400 bytes that
 * every GPU loading the library holds until exit: ncols > B200_NOALLOC_MAX_COLS is rejected with
 * B200_ERR_ARG (rows are not limited).  All calls on a device share that array: calls that overlap on the
 * GPU (different streams or threads) race -- serialise them.  Single process: a multi-rank caller sums
 * out[] over the ranks itself. */
#define B200_NOALLOC_MAX_COLS 32
int b200_dot_noalloc(const double* x, int64_t ldx, const double* y, int64_t ldy, int64_t n, int ncols,
                     double* out, void* stream);

/* ----------------------------------------------------------------- operator
 * Replaces Epetra_CrsMatrix::Multiply(false, X, Y)
 * (epetra/src/Epetra_CrsMatrix.cpp:3107-3285; kernels GeneralMV :3315-3413,
 *  GeneralMM :3486-3585) together with the Epetra_Import halo exchange
 * (epetra/src/Epetra_DistObject.cpp:226-291, Epetra_MultiVector.cpp:619-913)
 * and the column-map construction of Epetra_CrsGraph::MakeColMap
 * (epetra/src/Epetra_CrsGraph.cpp:1348-1524).
 *
 * Rows [row_begin,row_end) of an n_global x n_global matrix are given in CSR
 * with GLOBAL column ids; the row-block partition is the uniform linear map of
 * Epetra_Map(n,0,comm) (epetra/src/Epetra_BlockMap.cpp:83-90).  The library
 * builds the local column map (owned columns first, then ghosts grouped by
 * owner rank, ascending GID), the send/recv lists and the interior/boundary
 * row split.  Arrays are host pointers and are copied. */
int b200_op_create_csr(b200_ctx* ctx, int64_t n_global, int64_t row_begin, int64_t row_end,
                       const int32_t* rowptr, const int64_t* colind_global,
                       const double* values, b200_op** out);
int b200_op_destroy(b200_op* op);
/* y <- A x (trans==0) ; trans!=0 returns B200_ERR_UNSUPPORTED (GMRES never asks;
 * belos/src/BelosOperator.hpp:141 HasApplyTranspose() == false). x and y must
 * not alias (Epetra handles X==Y with a temp copy, Epetra_CrsMatrix.cpp:3152;
 * Belos never does it, belos/src/BelosOperatorTraits.hpp). */
int b200_op_apply(const b200_op* op, const b200_mv* x, b200_mv* y, int trans);
int b200_op_info(const b200_op* op, int64_t* n_local, int64_t* n_global, int64_t* nnz_local,
                 int64_t* n_ghost, int64_t* n_boundary_rows);

/* -------------------------------------------------------- host-side planning
 * Pure host functions (no device needed) — used by b200_op_create_csr and by
 * the CPU (gloo) tests of the multi-rank path. */
/* Epetra_Map(n, 0, comm) linear partition  epetra/src/Epetra_BlockMap.cpp:83-90 */
int b200_linear_partition(int64_t n_global, int nranks, int rank, int64_t* row_begin,
                          int64_t* row_end);
/* MakeColMap: returns number of ghosts; ghost_gids (capacity nnz) filled in local
 * order, colind_local filled with local ids, ghost_owner per ghost.
 * epetra/src/Epetra_CrsGraph.cpp:1426-1473 */
int64_t b200_build_colmap(int64_t n_global, int nranks, int rank, int64_t nnz,
                          const int64_t* colind_global, int32_t* colind_local,
                          int64_t* ghost_gids, int32_t* ghost_owner);

/* Halo plan, host part (replaces the Epetra_Import constructor, epetra/src/Epetra_Import.cpp:319-460):
 * recv side: per-owner counts and offsets into the ghost tail (ghosts are grouped by owner);
 * send side: global ids the peers asked for -> local row ids to pack (ExportLIDs). */
int b200_halo_recv_plan(int nranks, int64_t n_ghost, const int32_t* ghost_owner, int64_t* need_cnt,
                        int64_t* recv_off);
int b200_halo_send_index(int64_t n_global, int nranks, int rank, int64_t n_asked,
                         const int64_t* asked_gids, int32_t* send_idx);

/* Host-only: the tile plan the window-staged SpMM kernel (trilinos_b200/csrc/spmm_win.cu) streams, for a local
 * CSR block with local column ids (>= n_local: ghost columns) -- exported so that the CPU tests can replay
 * the kernel's arithmetic from the plan and compare it with Epetra_CrsMatrix::GeneralMM's row sums
 * (epetra/src/Epetra_CrsMatrix.cpp:3516-3534) bit for bit.  rows_per_tile: 128 | 256.  Two-call protocol:
 * tiles_out / blob_out NULL -> sizes only (ntiles x 128-byte descriptors, blob_bytes).  Returns 0, or 1 if the
 * format does not apply to this matrix (more than 8 x-runs per tile, rows longer than 255, > 25 % padding). */
int b200_spmm_window_plan(int64_t n_local, int64_t n_ghost, const int32_t* rowptr, const int32_t* colind_local,
                          const double* values, int rows_per_tile, int64_t* ntiles, int64_t* blob_bytes,
                          int32_t* blob_cap, int32_t* wcap, void* tiles_out, void* blob_out);

/* Galeri-style generators (galeri/src-epetra/Galeri_CrsMatrices.cpp:273-416,
 * CrsMatrices/Galeri_Cross2D.h, Galeri_Cross3D.h, Galeri_Recirc2D.h;
 * Brick3D: galeri/src-xpetra/Galeri_XpetraMatrixTypes.hpp:531-633) on the
 * "Linear" map.  name in {"Laplace2D","Laplace3D","Recirc2D","Brick3D","Tridiag",
 * "Diag"}; params: conv, diff (Recirc2D), a (Diag value / stencil centre override, 0 = default).
 * Two-call protocol: rowptr==NULL -> only *nnz is set. */
int b200_galeri_generate(const char* name, int nx, int ny, int nz, double conv, double diff,
                         int64_t row_begin, int64_t row_end, int64_t* nnz, int32_t* rowptr,
                         int64_t* colind_global, double* values);

/* ---- matrix ingest ------------------------------------------------
 * Harwell-Boeing file -> the rows [row_begin, row_end) (row_end < 0: all) of the matrix the reference
 * builds from it: Belos::Util::createEpetraProblem (belos/epetra/src/BelosEpetraUtils.cpp:81-200)
 * over Trilinos_Util_read_hb (triutils/src/Trilinos_Util_read_hb.cpp:49-230) -- real assembled square
 * matrices, symmetric storage mirrored, an explicit diagonal entry in every row (MSR), columns
 * sorted.  rhs / xexact (length row_end - row_begin) are filled when the file carries a full
 * right-hand side with an exact solution (*has_rhs = 1).  Two-call protocol: rowptr == NULL -> only
 * *n, *nnz, *has_rhs are set.  Output is what b200_op_create_csr takes. */
int b200_read_hb(const char* path, int64_t row_begin, int64_t row_end, int64_t* n, int64_t* nnz,
                 int32_t* rowptr, int64_t* colind_global, double* values, double* rhs, double* xexact,
                 int* has_rhs);
/* MatrixMarket "matrix coordinate real general|symmetric" (triutils/src/
 * Trilinos_Util_ReadMatrixMarket2Epetra.cpp): repeated entries are summed, no diagonal is added. */
int b200_read_matrix_market(const char* path, int64_t row_begin, int64_t row_end, int64_t* n,
                            int64_t* nnz, int32_t* rowptr, int64_t* colind_global, double* values);

/* ---- right preconditioning through the same boundary -------------
 * Element-wise product with a broadcast vector: mv(:,j) = scalar_this*mv(:,j) + scalar_ab * A(:,0) .* B(:,j)
 * Epetra_MultiVector::Multiply(double ScalarAB, A, B, double ScalarThis) with a single-vector A
 * (epetra/src/Epetra_MultiVector.cpp; scalar_this == 0 never reads mv).  B may alias mv. */
int b200_mv_multiply_elementwise(b200_mv* mv, double scalar_ab, const b200_mv* A, const b200_mv* B,
                                 double scalar_this);
/* Epetra_CrsMatrix::ExtractDiagonalCopy: a new one-column multivector on the operator's row map. */
int b200_op_extract_diagonal(const b200_op* op, b200_mv** diag);

/* Chebyshev polynomial smoother in D^-1 A, the reference's Ifpack_Chebyshev in point mode with a
 * zero starting solution (ifpack/src/Ifpack_Chebyshev.cpp: SetParameters :160-205, Compute
 * :283-372, ApplyInverse :466-725, PowerMethod :729-759).  Fields follow the ParameterList keys
 * "chebyshev: degree", ": max eigenvalue" (< 0: estimated by the power method with eig_iters
 * iterations from a start vector drawn with `seed`), ": ratio eigenvalue", ": min diagonal
 * value", ": eigenvalue max iterations". */
typedef struct b200_cheby b200_cheby;
typedef struct {
  int degree;
  double lambda_max, eig_ratio, min_diag;
  int eig_iters;
  unsigned int seed;
} b200_cheby_params;
void b200_cheby_default_params(b200_cheby_params* p); /* 1, -1, 30, 0, 10, 1 (Ifpack defaults) */
/* constructor + SetParameters + Initialize + Compute; the operator must outlive the smoother */
int b200_cheby_create(b200_op* op, const b200_cheby_params* p, b200_cheby** out);
/* ApplyInverse: y = p(D^-1 A) D^-1 x; x and y are n x k multivectors of the operator's map */
int b200_cheby_apply(b200_cheby* c, const b200_mv* x, b200_mv* y);
int b200_cheby_lambda(const b200_cheby* c, double* lambda_max, double* lambda_min);
int b200_cheby_destroy(b200_cheby* c);

#ifdef __cplusplus
}
#endif
#endif /* B200_BELOS_H */
