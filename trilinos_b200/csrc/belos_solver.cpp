// libb200belos_solver.so: unmodified Belos::BlockGmresSolMgr on the B200 plugin classes.
// See include/b200_belos_solver.h for the reference code each entry point replaces.
#include "b200_belos_solver.h"

#include <chrono>
#include <cstring>
#include <iostream>
#include <sstream>

#include "B200BelosAdapter.hpp"
#include "B200ResidualHistory.hpp"
#include "BelosBlockCGSolMgr.hpp"
#include "BelosBlockGmresSolMgr.hpp"
#include "BelosPseudoBlockCGSolMgr.hpp"
#include "BelosPseudoBlockGmresSolMgr.hpp"
#include "BelosLinearProblem.hpp"
#include "BelosMVOPTester.hpp"
#include "BelosOrthoManagerFactory.hpp"
#include "BelosOutputManager.hpp"
#include "Teuchos_ParameterList.hpp"
#include "Teuchos_RCP.hpp"
#include "Teuchos_TimeMonitor.hpp"

namespace b200 {
int fail(int code, const char* msg);
}

typedef double ST;
typedef Belos::MultiVec<ST> MV;
typedef Belos::Operator<ST> OP;
typedef Belos::MultiVecTraits<ST, MV> MVT;
typedef Belos::OperatorTraits<ST, MV, OP> OPT;

extern "C" void b200_gmres_default_params(b200_gmres_params* p) {
  if (!p) return;
  p->block_size = 1;
  p->num_blocks = 300;
  p->max_iters = 1000;
  p->max_restarts = 20;
  p->tol = 1e-8;
  p->ortho = "DGKS";
  p->verbosity = 0;
  p->output_frequency = -1;
  p->adaptive_block_size = 1;
  p->flexible = 0;
  p->show_timers = 0;
}

extern "C" int b200_belos_block_gmres_solve(b200_op* A_h, b200_mv* X_h, b200_mv* B_h,
                                            const b200_gmres_params* p, b200_gmres_result* r) {
  return b200_belos_solve("BlockGmres", A_h, X_h, B_h, p, r);
}

extern "C" int b200_belos_solve(const char* solver_name, b200_op* A_h, b200_mv* X_h, b200_mv* B_h,
                                const b200_gmres_params* p, b200_gmres_result* r) {
  return b200_belos_solve_prec(solver_name, A_h, nullptr, 0, X_h, B_h, p, r);
}

extern "C" int b200_belos_solve_prec(const char* solver_name, b200_op* A_h, b200_cheby* prec,
                                     int left, b200_mv* X_h, b200_mv* B_h,
                                     const b200_gmres_params* p, b200_gmres_result* r) {
  return b200_belos_solve_hist(solver_name, A_h, prec, left, X_h, B_h, p, r, nullptr, 0, nullptr);
}

extern "C" int b200_belos_solve_hist(const char* solver_name, b200_op* A_h, b200_cheby* prec,
                                     int left, b200_mv* X_h, b200_mv* B_h,
                                     const b200_gmres_params* p, b200_gmres_result* r,
                                     double* hist, int hist_cap_rows, int* hist_rows) {
  using Teuchos::RCP;
  using Teuchos::rcp;
  if (!solver_name || !A_h || !X_h || !B_h || !p || !r)
    return b200::fail(B200_ERR_ARG, "solve: null argument");
  const std::string which(solver_name);
  if (which != "BlockGmres" && which != "PseudoBlockGmres" && which != "BlockCG" &&
      which != "PseudoBlockCG")
    return b200::fail(B200_ERR_ARG,
                      "solve: unknown solver manager (BlockGmres | PseudoBlockGmres | BlockCG | PseudoBlockCG)");
  std::memset(r, 0, sizeof(*r));
  try {
    // Non-owning views of the caller's handles: CloneView over all columns shares storage.
    const int numrhs = b200_mv_num_vecs(B_h);
    std::vector<int> all(numrhs);
    for (int i = 0; i < numrhs; ++i) all[i] = i;
    b200_mv *xv = nullptr, *bv = nullptr;
    if (b200_mv_clone_view(X_h, all.data(), numrhs, &xv) != 0) return B200_ERR_ARG;
    if (b200_mv_clone_view(B_h, all.data(), numrhs, &bv) != 0) {
      b200_mv_destroy(xv);
      return B200_ERR_ARG;
    }
    RCP<MV> X = rcp(new Belos::B200MultiVec(xv));
    RCP<MV> B = rcp(new Belos::B200MultiVec(bv));
    RCP<OP> A = rcp(new Belos::B200CrsOp(A_h, false));

    const ptrdiff_t n_global = MVT::GetGlobalLength(*B);
    int maxiters = p->max_iters;
    if (maxiters == -1) maxiters = (int)(n_global / std::max(1, p->block_size)) - 1;

    Teuchos::ParameterList belosList;
    belosList.set("Num Blocks", p->num_blocks);
    belosList.set("Block Size", p->block_size);
    belosList.set("Maximum Iterations", maxiters);
    belosList.set("Maximum Restarts", p->max_restarts);
    belosList.set("Convergence Tolerance", p->tol);
    belosList.set("Orthogonalization", std::string(p->ortho ? p->ortho : "DGKS"));
    belosList.set("Adaptive Block Size", p->adaptive_block_size != 0);
    if (p->flexible) belosList.set("Flexible Gmres", true);
    int verbosity = Belos::Errors + Belos::Warnings + p->verbosity;
    if (p->show_timers) verbosity |= Belos::TimingDetails;
    belosList.set("Verbosity", verbosity);
    // Belos' own messages (warnings, status-test output, timers) go to stderr: stdout stays the caller's
    belosList.set("Output Stream", Teuchos::rcp(&std::cerr, false));
    if (p->output_frequency > 0) belosList.set("Output Frequency", p->output_frequency);

    Belos::LinearProblem<ST, MV, OP> problem(A, X, B);
    if (prec) {  // BelosLinearProblem.hpp:1057-1071 (right) / :1040-1056 (left)
      RCP<OP> M = rcp(new Belos::B200ChebyshevOp(prec, false));
      if (left) problem.setLeftPrec(M); else problem.setRightPrec(M);
    }
    if (!problem.setProblem())
      return b200::fail(B200_ERR_ARG, "Belos::LinearProblem failed to set up correctly");

    Teuchos::TimeMonitor::zeroOutTimers();
    // every manager is the reference's own class template, instantiated on the plugin classes
    RCP<Belos::SolverManager<ST, MV, OP> > solver;
    RCP<Belos::LinearProblem<ST, MV, OP> > prob = rcp(&problem, false);
    RCP<Teuchos::ParameterList> pl = rcp(&belosList, false);
    RCP<Belos::B200ResidualHistory<ST, MV, OP> > rec;
    if (hist_rows) *hist_rows = 0;
    if (which == "BlockGmres") {
      RCP<Belos::BlockGmresSolMgr<ST, MV, OP> > bg = rcp(new Belos::BlockGmresSolMgr<ST, MV, OP>(prob, pl));
      if (hist_cap_rows > 0 && hist) {
        // setDebugStatusTest + setProblem: solve() rebuilds its status-test combo with the recorder in it
        rec = rcp(new Belos::B200ResidualHistory<ST, MV, OP>());
        bg->setDebugStatusTest(rec);
        bg->setProblem(prob);
      }
      solver = bg;
    } else if (which == "PseudoBlockGmres") {
      belosList.remove("Adaptive Block Size");
      belosList.remove("Flexible Gmres", false);
      solver = rcp(new Belos::PseudoBlockGmresSolMgr<ST, MV, OP>(prob, pl));
    } else {
      belosList.remove("Num Blocks");
      belosList.remove("Maximum Restarts");
      belosList.remove("Flexible Gmres", false);
      if (which == "BlockCG") {
        solver = rcp(new Belos::BlockCGSolMgr<ST, MV, OP>(prob, pl));
      } else {
        belosList.remove("Block Size");
        belosList.remove("Adaptive Block Size");
        belosList.remove("Orthogonalization");
        solver = rcp(new Belos::PseudoBlockCGSolMgr<ST, MV, OP>(prob, pl));
      }
    }
    b200_ctx_sync(b200_mv_ctx(B_h));
    const auto t0 = std::chrono::steady_clock::now();
    const Belos::ReturnType ret = solver->solve();
    b200_ctx_sync(b200_mv_ctx(B_h));
    const auto t1 = std::chrono::steady_clock::now();

    r->converged = (ret == Belos::Converged) ? 1 : 0;
    r->num_iters = solver->getNumIters();
    r->achieved_tol = solver->achievedTol();
    r->loss_of_accuracy = solver->isLOADetected() ? 1 : 0;
    r->solve_seconds = std::chrono::duration<double>(t1 - t0).count();
    r->num_rhs = numrhs;
    if (nonnull(rec)) {
      const int rows = rec->copy_out(hist, hist_cap_rows, numrhs);
      if (hist_rows) *hist_rows = rows;
    }

    // true residual check, as the reference driver does after solve()
    std::vector<double> actual(numrhs), rhs(numrhs);
    RCP<MV> resid = MVT::Clone(*B, numrhs);
    OPT::Apply(*A, *X, *resid);
    MVT::MvAddMv(-1.0, *resid, 1.0, *B, *resid);
    MVT::MvNorm(*resid, actual);
    MVT::MvNorm(*B, rhs);
    for (int i = 0; i < numrhs && i < 64; ++i)
      r->true_rel_resid[i] = rhs[i] > 0 ? actual[i] / rhs[i] : actual[i];
    return 0;
  } catch (const Belos::B200MultiVecFailure& e) {
    // a C-ABI call returned non-zero: hand its own status (CUDA, NCCL, timeout, shape ...) back
    const int st = b200_last_status();
    return b200::fail(st != 0 ? st : B200_ERR_ARG, e.what());
  } catch (const Belos::B200OpFailure& e) {
    const int st = b200_last_status();
    return b200::fail(st != 0 ? st : B200_ERR_ARG, e.what());
  } catch (const Belos::BelosError& e) {
    // Belos' own failures (GmresIterationOrthoFailure rethrown by the manager, LinearProblemError,
    // StatusTestError ...): BelosBlockGmresSolMgr.hpp:1161-1187
    return b200::fail(B200_ERR_BELOS, e.what());
  } catch (const std::invalid_argument& e) {
    return b200::fail(B200_ERR_ARG, e.what());
  } catch (const std::exception& e) {
    return b200::fail(B200_ERR_BELOS, e.what());
  }
}

extern "C" int b200_belos_timer_summary(char* buf, int buflen) {
  if (!buf || buflen <= 0) return B200_ERR_ARG;
  std::ostringstream os;
  Teuchos::TimeMonitor::summarize(os);
  const std::string s = os.str();
  std::strncpy(buf, s.c_str(), (size_t)buflen - 1);
  buf[buflen - 1] = 0;
  return 0;
}

extern "C" int b200_belos_mvop_test(b200_ctx* ctx, int n, int ncols, int verbose) {
  using Teuchos::RCP;
  using Teuchos::rcp;
  if (!ctx || n <= 0 || ncols <= 0) return b200::fail(B200_ERR_ARG, "mvop_test: bad arguments");
  try {
    RCP<Belos::OutputManager<ST> > om = rcp(new Belos::OutputManager<ST>());
    if (verbose) om->setVerbosity(Belos::Warnings + Belos::Errors + Belos::Debug);
    // tridiagonal operator, as belos/epetra/test/MVOPTester/cxx_main.cpp:112-148
    int rank = 0, nranks = 1;
    b200_ctx_rank(ctx, &rank, &nranks);
    int64_t rb, re;
    b200_linear_partition(n, nranks, rank, &rb, &re);
    int64_t nnz = 0;
    b200_galeri_generate("Tridiag", n, 1, 1, 0, 0, rb, re, &nnz, nullptr, nullptr, nullptr);
    std::vector<int32_t> rp(re - rb + 1);
    std::vector<int64_t> ci(nnz);
    std::vector<double> va(nnz);
    b200_galeri_generate("Tridiag", n, 1, 1, 0, 0, rb, re, &nnz, rp.data(), ci.data(), va.data());
    b200_op* op_h = nullptr;
    if (b200_op_create_csr(ctx, n, rb, re, rp.data(), ci.data(), va.data(), &op_h) != 0) return B200_ERR_ARG;
    RCP<OP> op = rcp(new Belos::B200CrsOp(op_h, true));
    RCP<MV> ivec = rcp(new Belos::B200MultiVec(ctx, re - rb, n, rb, ncols));
    MVT::MvRandom(*ivec);
    bool ok = Belos::TestMultiVecTraits<ST, MV>(om, ivec);
    if (!ok) return b200::fail(1, "Belos::TestMultiVecTraits() returned false");
    ok = Belos::TestOperatorTraits<ST, MV, OP>(om, ivec, op);
    if (!ok) return b200::fail(2, "Belos::TestOperatorTraits() returned false");
    return 0;
  } catch (const std::exception& e) {
    return b200::fail(B200_ERR_ARG, e.what());
  }
}

namespace {
typedef Teuchos::SerialDenseMatrix<int, ST> DM;

double fro(const DM& M) { return M.normFrobenius(); }

// || X^T Y ||_F  and  || X^T X - I ||_F, as OrthoManager::orthogError / orthonormError measure them
double orthog_err(const MV& X, const MV& Y) {
  DM Z(MVT::GetNumberVecs(X), MVT::GetNumberVecs(Y));
  MVT::MvTransMv(1.0, X, Y, Z);
  return fro(Z);
}
double orthonorm_err(const MV& X) {
  const int k = MVT::GetNumberVecs(X);
  DM Z(k, k);
  MVT::MvTransMv(1.0, X, X, Z);
  for (int i = 0; i < k; ++i) Z(i, i) -= 1.0;
  return fro(Z);
}
double max_col_norm(const MV& X) {
  std::vector<double> nrm(MVT::GetNumberVecs(X));
  MVT::MvNorm(X, nrm);
  double m = 0.0;
  for (double v : nrm) m = std::max(m, v);
  return m;
}
}  // namespace

// The checks of Belos::Test::OrthoManagerTester (belos/src/BelosOrthoManagerTest.hpp:353-760, testNormalize
// :1135-1298, testProjectAndNormalizeNew :1300-1475), restated: that header does not compile against
// Belos::MultiVec<double> in this reference version (it calls MultiVecTraits::HasConstantStride and a misspelt
// Teuchos::randomSyncedmatrix), so the same quantities are measured here with the same bounds:
//   normalize(S, B):                 rank, || S^T S - I ||, || S_in - S B || / || S_in ||
//   projectAndNormalize(S, C, B, Q): the same plus || Q_i^T S || and || S_in - sum Q_i C_i - S B || / || S_in ||
// on a random block and, for rank-revealing managers, on a block with a duplicated column and a zero column.
extern "C" int b200_belos_ortho_test(b200_ctx* ctx, const char* ortho, int n, int ncols, int size_x1,
                                     int size_x2, int rank_revealing, int verbose, int* num_failed) {
  using Teuchos::RCP;
  using Teuchos::rcp;
  if (!ctx || !ortho || !num_failed || n <= 0 || ncols <= 0 || size_x1 <= 0 || size_x2 <= 0)
    return b200::fail(B200_ERR_ARG, "ortho_test: bad arguments");
  *num_failed = -1;
  try {
    RCP<Belos::OutputManager<ST> > om = rcp(new Belos::OutputManager<ST>(
        Belos::Errors | (verbose ? (Belos::Warnings | Belos::Debug) : 0), rcp(&std::cerr, false)));
    int rank = 0, nranks = 1;
    b200_ctx_rank(ctx, &rank, &nranks);
    int64_t rb, re;
    b200_linear_partition(n, nranks, rank, &rb, &re);
    RCP<MV> S = rcp(new Belos::B200MultiVec(ctx, re - rb, n, rb, ncols));
    MVT::MvRandom(*S);
    Belos::OrthoManagerFactory<ST, MV, OP> factory;
    RCP<Belos::OrthoManager<ST, MV> > OM =
        factory.makeOrthoManager(std::string(ortho), Teuchos::null, om, "Belos", Teuchos::null);
    const double tol = 1e-12 * std::sqrt((double)n);  // ~ the tester's 100 * eps * sqrt(n) scale
    int failed = 0;
    std::ostream& os = std::cerr;
    auto check = [&](const char* what, double got, double bound) {
      const bool ok = got <= bound;  // (false for NaN)
      if (!ok) ++failed;
      if (verbose || !ok) os << "ortho_test[" << ortho << "] " << what << ": " << got << (ok ? " <= " : " > ") << bound << std::endl;
    };
    // orthonormal bases X1, X2 to project against (built with the manager under test, as the tester does)
    RCP<MV> X1 = MVT::Clone(*S, size_x1), X2 = MVT::Clone(*S, size_x2);
    MVT::MvRandom(*X1);
    MVT::MvRandom(*X2);
    {
      int r1 = OM->normalize(*X1, Teuchos::null);
      check("normalize(X1): rank deficit", (double)(size_x1 - r1), 0.0);
      check("normalize(X1): || X1^T X1 - I ||", orthonorm_err(*X1), tol);
      Teuchos::Array<RCP<DM> > C(1, Teuchos::null);
      Teuchos::Array<RCP<const MV> > Q(1, X1);
      int r2 = OM->projectAndNormalize(*X2, C, Teuchos::null, Q);
      check("projectAndNormalize(X2 | X1): rank deficit", (double)(size_x2 - r2), 0.0);
      check("|| X2^T X2 - I ||", orthonorm_err(*X2), tol);
      check("|| X1^T X2 ||", orthog_err(*X1, *X2), tol);
    }
    for (int variant = 0; variant < (rank_revealing ? 2 : 1); ++variant) {
      // variant 0: random full-rank block; variant 1: column 1 := column 0 (dependent) and the last column := 0
      RCP<MV> Sin = MVT::CloneCopy(*S);
      int expect_rank = ncols;
      if (variant == 1 && ncols >= 3) {
        std::vector<int> c0(1, 0), c1(1, 1), cl(1, ncols - 1);
        RCP<const MV> s0 = MVT::CloneView(*Sin, c0);
        MVT::SetBlock(*s0, c1, *Sin);
        RCP<MV> sl = MVT::CloneViewNonConst(*Sin, cl);
        MVT::MvInit(*sl, 0.0);
        expect_rank = ncols - 2;
      }
      const double scale = std::max(max_col_norm(*Sin), 1e-300);
      {  // normalize
        RCP<MV> W = MVT::CloneCopy(*Sin);
        RCP<DM> B = rcp(new DM(ncols, ncols));
        const int r = OM->normalize(*W, B);
        check(variant ? "normalize(deficient): |rank - expected|" : "normalize: |rank - expected|",
              std::fabs((double)(r - expect_rank)), 0.0);
        if (r == ncols) check("normalize: || S^T S - I ||", orthonorm_err(*W), tol);
        else {
          std::vector<int> lead(std::max(r, 1));
          for (int i = 0; i < (int)lead.size(); ++i) lead[i] = i;
          RCP<const MV> Wl = MVT::CloneView(*W, lead);
          check("normalize(deficient): || S_r^T S_r - I ||", orthonorm_err(*Wl), tol);
        }
        RCP<MV> Rs = MVT::CloneCopy(*Sin);
        MVT::MvTimesMatAddMv(-1.0, *W, *B, 1.0, *Rs);
        check("normalize: || S_in - S B || / || S_in ||", max_col_norm(*Rs) / scale, 100 * tol);
      }
      {  // projectAndNormalize against {X1, X2}
        RCP<MV> W = MVT::CloneCopy(*Sin);
        RCP<DM> B = rcp(new DM(ncols, ncols));
        Teuchos::Array<RCP<DM> > C(2);
        C[0] = rcp(new DM(size_x1, ncols));
        C[1] = rcp(new DM(size_x2, ncols));
        Teuchos::Array<RCP<const MV> > Q(2);
        Q[0] = X1;
        Q[1] = X2;
        const int r = OM->projectAndNormalize(*W, C, B, Q);
        check("projectAndNormalize: |rank - expected|", std::fabs((double)(r - expect_rank)), 0.0);
        std::vector<int> lead(std::max(r, 1));
        for (int i = 0; i < (int)lead.size(); ++i) lead[i] = i;
        RCP<const MV> Wl = MVT::CloneView(*W, lead);
        check("projectAndNormalize: || S^T S - I ||", orthonorm_err(*Wl), tol);
        check("projectAndNormalize: || X1^T S ||", orthog_err(*X1, *Wl), tol);
        check("projectAndNormalize: || X2^T S ||", orthog_err(*X2, *Wl), tol);
        RCP<MV> Rs = MVT::CloneCopy(*Sin);
        MVT::MvTimesMatAddMv(-1.0, *X1, *C[0], 1.0, *Rs);
        MVT::MvTimesMatAddMv(-1.0, *X2, *C[1], 1.0, *Rs);
        MVT::MvTimesMatAddMv(-1.0, *W, *B, 1.0, *Rs);
        check("projectAndNormalize: || S_in - X1 C1 - X2 C2 - S B || / || S_in ||", max_col_norm(*Rs) / scale, 100 * tol);
      }
    }
    *num_failed = failed;
    return 0;
  } catch (const Belos::B200MultiVecFailure& e) {
    const int st = b200_last_status();
    return b200::fail(st != 0 ? st : B200_ERR_ARG, e.what());
  } catch (const std::exception& e) {
    return b200::fail(B200_ERR_BELOS, e.what());
  }
}
