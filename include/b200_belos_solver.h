/*
 * b200_belos_solver.h — C entry points of libb200belos_solver.so: the UNMODIFIED
 * Belos::BlockGmresSolMgr<double, Belos::MultiVec<double>, Belos::Operator<double>>
 * (belos/src/BelosBlockGmresSolMgr.hpp:916-1303, compiled from the reference headers where
 * they lie) driving the B200 plugin classes of include/B200BelosAdapter.hpp.
 *
 * b200_belos_block_gmres_solve replaces the body of the reference driver
 * belos/epetra/example/BlockGmres/BlockGmresEpetraGaleriEx.cpp:170-262 (ParameterList,
 * LinearProblem::setProblem, BlockGmresSolMgr::solve, true-residual check);
 * b200_belos_mvop_test replaces belos/epetra/test/MVOPTester/cxx_main.cpp:91-176
 * (Belos::TestMultiVecTraits / TestOperatorTraits, belos/src/BelosMVOPTester.hpp).
 */
#ifndef B200_BELOS_SOLVER_H
#define B200_BELOS_SOLVER_H

#include "b200_belos.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
  int block_size;          /* "Block Size"            (default 1)   */
  int num_blocks;          /* "Num Blocks"            (default 300) */
  int max_iters;           /* "Maximum Iterations"    (-1: n/block_size - 1 as the example driver) */
  int max_restarts;        /* "Maximum Restarts"      (default 20)  */
  double tol;              /* "Convergence Tolerance" (default 1e-8) */
  const char* ortho;       /* "Orthogonalization": "DGKS" | "ICGS" | "IMGS" */
  int verbosity;           /* Belos::MsgType bitmask; 0 = Errors only */
  int output_frequency;    /* "Output Frequency" (-1: never) */
  int adaptive_block_size; /* "Adaptive Block Size" (default 1) */
  int flexible;            /* "Flexible Gmres" (default 0) */
  int show_timers;         /* print the Teuchos::TimeMonitor summary after the solve */
} b200_gmres_params;

typedef struct {
  int converged;           /* Belos::Converged == 1 */
  int num_iters;           /* BlockGmresSolMgr::getNumIters() */
  double achieved_tol;     /* BlockGmresSolMgr::achievedTol() */
  int loss_of_accuracy;    /* BlockGmresSolMgr::isLOADetected() */
  double solve_seconds;    /* wall time of solve() (host clock) */
  int num_rhs;
  double true_rel_resid[64]; /* ||B - A X||_2 / ||B||_2 per right-hand side (first 64) */
} b200_gmres_result;

void b200_gmres_default_params(b200_gmres_params* p);

/* X (initial guess in, solution out) and B are n x numrhs multivectors of the operator's map. */
int b200_belos_block_gmres_solve(b200_op* A, b200_mv* X, b200_mv* B, const b200_gmres_params* p,
                                 b200_gmres_result* r);

/* The same driver with the solver manager chosen by name -- every one is the reference's own class
 * template instantiated on <double, Belos::MultiVec<double>, Belos::Operator<double>>, so all of
 * them run on the same kernels:
 *   "BlockGmres"        belos/src/BelosBlockGmresSolMgr.hpp
 *   "PseudoBlockGmres"  belos/src/BelosPseudoBlockGmresSolMgr.hpp  (block_size = number of systems
 *                       advanced together; adaptive_block_size / flexible ignored)
 *   "BlockCG"           belos/src/BelosBlockCGSolMgr.hpp           (num_blocks / max_restarts ignored)
 *   "PseudoBlockCG"     belos/src/BelosPseudoBlockCGSolMgr.hpp     (also ignores block_size / ortho)
 * Reference drivers: belos/epetra/test/BlockGmres/test_pseudo_bl_gmres_hb.cpp,
 * belos/epetra/test/BlockCG/test_bl_cg_hb.cpp, test_pseudo_bl_cg_hb.cpp. */
int b200_belos_solve(const char* solver, b200_op* A, b200_mv* X, b200_mv* B,
                     const b200_gmres_params* p, b200_gmres_result* r);

/* The same with a Chebyshev smoother (b200_cheby, include/b200_belos.h) as preconditioner:
 * LinearProblem::setRightPrec (left == 0; belos/src/BelosLinearProblem.hpp:1057-1071) or setLeftPrec,
 * wrapped as Belos::B200ChebyshevOp the way the reference wraps Ifpack objects in
 * Belos::EpetraPrecOp (belos/epetra/src/BelosEpetraAdapter.hpp:280-438).  prec == NULL: none. */
int b200_belos_solve_prec(const char* solver, b200_op* A, b200_cheby* prec, int left, b200_mv* X,
                          b200_mv* B, const b200_gmres_params* p, b200_gmres_result* r);

/* b200_belos_solve_prec that also records, for solver "BlockGmres", the native residual norms Belos
 * holds before every block iteration (Iteration::getNativeResiduals, belos/src/BelosBlockGmresIter.hpp:542-557)
 * through a status test attached with BlockGmresSolMgr::setDebugStatusTest (BelosBlockGmresSolMgr.hpp:240):
 * hist[row * numrhs + j], row = block iteration; *hist_rows = rows recorded (may exceed hist_cap_rows,
 * only the first hist_cap_rows are stored).  Host data only: no extra kernel, same iteration.
 *
 * Status of the three solve entry points: 0, or the status of the C-ABI call that failed underneath
 * (B200_ERR_CUDA / _NCCL / _TIMEOUT / _SHAPE ...), B200_ERR_BELOS for exceptions raised by Belos itself,
 * B200_ERR_ARG for bad arguments / parameters. */
int b200_belos_solve_hist(const char* solver, b200_op* A, b200_cheby* prec, int left, b200_mv* X,
                          b200_mv* B, const b200_gmres_params* p, b200_gmres_result* r,
                          double* hist, int hist_cap_rows, int* hist_rows);

/* Runs Belos::TestMultiVecTraits and Belos::TestOperatorTraits (the reference's contract test)
 * on a B200MultiVec of n rows x ncols columns and a tridiagonal B200CrsOp. 0 = passed. */
int b200_belos_mvop_test(b200_ctx* ctx, int n, int ncols, int verbose);

/* The checks of the reference's orthogonalisation-manager test, Belos::Test::OrthoManagerTester::runTests
 * (belos/src/BelosOrthoManagerTest.hpp:353-760: normalize / projectAndNormalize on random and -- when
 * rank_revealing != 0 -- rank-deficient blocks; orthonormality, orthogonality against the projected bases and
 * reconstruction of the input), restated in belos_solver.cpp (that header does not compile against
 * Belos::MultiVec<double> in this reference version), on
 * B200MultiVec blocks of n rows with the manager OrthoManagerFactory makes for `ortho`
 * ("TSQR" | "DGKS" | "ICGS" | "IMGS" | "Simple"), as belos/epetra/test/OrthoManager/belos_orthomanager_epetra.cpp
 * does for Epetra.  "TSQR" = Belos::TsqrOrthoManager on the factorExplicit / revealRank hooks of the plugin class
 * (SURVEY 8 f.3).  *num_failed = failed checks (0 = all passed). */
int b200_belos_ortho_test(b200_ctx* ctx, const char* ortho, int n, int ncols, int size_x1, int size_x2,
                          int rank_revealing, int verbose, int* num_failed);

/* Timer table of the last solve (Teuchos::TimeMonitor::summarize) written into buf. */
int b200_belos_timer_summary(char* buf, int buflen);

#ifdef __cplusplus
}
#endif
#endif
