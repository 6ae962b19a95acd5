// oracle/ref_harness.cpp — C entry points onto the REFERENCE's own compiled code
// (Epetra + Galeri + Belos/Epetra adapter + unmodified Belos::BlockGmresSolMgr), built by
// oracle/Makefile.ref from the sources where they lie under the reference tree into
// oracle/_ref/libref_epetra_belos.so.
//
// TEST INFRASTRUCTURE ONLY: used by tests/ (golden-vector generation and parity checks) and by
// bench.py's cpu_baseline / `--impl reference` legs.  The product never links it.
//
// Everything numerical below is executed by reference code:
//   ref_galeri            Galeri::CreateMap("Linear") + Galeri::CreateCrsMatrix      galeri/src-epetra/Galeri_{Maps,CrsMatrices}.cpp
//   ref_apply             Epetra_CrsMatrix::Multiply(false,X,Y)                       epetra/src/Epetra_CrsMatrix.cpp:3107
//   ref_mv_*              Belos::MultiVecTraits<double,Epetra_MultiVector>            belos/epetra/src/BelosEpetraAdapter.hpp:448-1064
//   ref_block_gmres_csr   Belos::BlockGmresSolMgr<double,Epetra_MultiVector,Epetra_Operator>, driver body as in
//                         belos/epetra/example/BlockGmres/BlockGmresEpetraGaleriEx.cpp:170-262
#include <chrono>
#include <cstdint>
#include <cstring>
#include <sstream>
#include <string>
#include <vector>

#include "BelosBlockCGSolMgr.hpp"
#include "BelosBlockGmresSolMgr.hpp"
#include "BelosPseudoBlockCGSolMgr.hpp"
#include "BelosPseudoBlockGmresSolMgr.hpp"
#include "BelosConfigDefs.hpp"
#include "BelosEpetraAdapter.hpp"
#include "BelosLinearProblem.hpp"
#include "B200ResidualHistory.hpp"
#include "Epetra_CrsMatrix.h"
#include "Epetra_Map.h"
#include "Epetra_MultiVector.h"
#include "Epetra_SerialComm.h"
#include "Epetra_Util.h"
#include "Epetra_Vector.h"
#include "Galeri_CrsMatrices.h"
#include "Galeri_Maps.h"
#include "Ifpack_Chebyshev.h"
#include "Trilinos_Util.h"
#include "Trilinos_Util_ReadMatrixMarket2Epetra.h"
#include "Teuchos_ParameterList.hpp"
#include "Teuchos_RCP.hpp"
#include "Teuchos_SerialDenseMatrix.hpp"
#include "Teuchos_TimeMonitor.hpp"

typedef double ST;
typedef Epetra_MultiVector MV;
typedef Epetra_Operator OP;
typedef Belos::MultiVecTraits<ST, MV> MVT;
typedef Belos::OperatorTraits<ST, MV, OP> OPT;
using Teuchos::RCP;
using Teuchos::rcp;

static thread_local std::string g_err;
static int fail(const std::exception& e) {
  g_err = e.what();
  return -1;
}

extern "C" const char* ref_last_error() { return g_err.c_str(); }

// layout-compatible with b200_gmres_params / b200_gmres_result (include/b200_belos_solver.h)
struct ref_gmres_params {
  int block_size, num_blocks, max_iters, max_restarts;
  double tol;
  const char* ortho;
  int verbosity, output_frequency, adaptive_block_size, flexible, show_timers;
};
struct ref_gmres_result {
  int converged, num_iters;
  double achieved_tol;
  int loss_of_accuracy;
  double solve_seconds;
  int num_rhs;
  double true_rel_resid[64];
};

namespace {

RCP<Epetra_CrsMatrix> csr_to_epetra(const Epetra_Map& map, int64_t n, const int32_t* rowptr,
                                    const int32_t* colind, const double* values) {
  std::vector<int> per_row((size_t)n);
  for (int64_t i = 0; i < n; ++i) per_row[i] = rowptr[i + 1] - rowptr[i];
  RCP<Epetra_CrsMatrix> A = rcp(new Epetra_CrsMatrix(Copy, map, per_row.data(), true));
  for (int64_t i = 0; i < n; ++i) {
    const int cnt = per_row[i];
    if (cnt > 0)
      A->InsertGlobalValues((int)i, cnt, const_cast<double*>(values + rowptr[i]),
                            const_cast<int*>(reinterpret_cast<const int*>(colind + rowptr[i])));
  }
  A->FillComplete();
  A->OptimizeStorage();
  return A;
}

// View-mode multivector over caller memory (no copy): Epetra_MultiVector(View, map, ptr, lda, k)
inline MV view(const Epetra_Map& map, const double* p, int64_t ld, int k) {
  return MV(View, map, const_cast<double*>(p), (int)ld, k);
}

}  // namespace

// Galeri matrix on the "Linear" map, serial comm.  Two-call protocol: rowptr == NULL -> only nnz.
extern "C" int ref_galeri(const char* name, int nx, int ny, int nz, double conv, double diff,
                          int64_t* nnz, int32_t* rowptr, int32_t* colind, double* values) {
  try {
    Epetra_SerialComm comm;
    Teuchos::ParameterList gl;
    const std::string nm(name);
    const bool is3d = (nm == "Laplace3D");
    gl.set("n", is3d ? nx * ny * nz : nx * ny);
    gl.set("nx", nx);
    gl.set("ny", ny);
    if (is3d) gl.set("nz", nz);
    if (nm == "Recirc2D") {
      gl.set("conv", conv);
      gl.set("diff", diff);
    }
    RCP<Epetra_Map> map = rcp(Galeri::CreateMap("Linear", comm, gl));
    RCP<Epetra_CrsMatrix> A = rcp(Galeri::CreateCrsMatrix(nm, &*map, gl));
    const int n = A->NumMyRows();
    *nnz = A->NumMyNonzeros();
    if (!rowptr) return 0;
    int64_t pos = 0;
    for (int i = 0; i < n; ++i) {
      int cnt;
      double* v;
      int* c;
      A->ExtractMyRowView(i, cnt, v, c);
      rowptr[i] = (int32_t)pos;
      for (int j = 0; j < cnt; ++j) {
        colind[pos] = A->GCID(c[j]);
        values[pos] = v[j];
        ++pos;
      }
    }
    rowptr[n] = (int32_t)pos;
    return 0;
  } catch (const std::exception& e) {
    return fail(e);
  }
}

extern "C" int ref_apply(int64_t n, const int32_t* rowptr, const int32_t* colind,
                         const double* values, const double* X, int64_t ldx, int k, double* Y,
                         int64_t ldy) {
  try {
    Epetra_SerialComm comm;
    Epetra_Map map((int)n, 0, comm);
    RCP<Epetra_CrsMatrix> A = csr_to_epetra(map, n, rowptr, colind, values);
    MV x = view(map, X, ldx, k), y = view(map, Y, ldy, k);
    OPT::Apply(*A, x, y);
    return 0;
  } catch (const std::exception& e) {
    return fail(e);
  }
}

extern "C" int ref_mv_trans_mv(int64_t n, int p, int q, double alpha, const double* A, int64_t lda,
                               const double* V, int64_t ldv, double* C, int ldc) {
  try {
    Epetra_SerialComm comm;
    Epetra_Map map((int)n, 0, comm);
    MV a = view(map, A, lda, p), v = view(map, V, ldv, q);
    Teuchos::SerialDenseMatrix<int, double> B(Teuchos::View, C, ldc, p, q);
    MVT::MvTransMv(alpha, a, v, B);
    return 0;
  } catch (const std::exception& e) {
    return fail(e);
  }
}

extern "C" int ref_mv_times_mat_add_mv(int64_t n, int p, int q, double alpha, const double* A,
                                       int64_t lda, const double* B, int ldb, double beta,
                                       double* V, int64_t ldv) {
  try {
    Epetra_SerialComm comm;
    Epetra_Map map((int)n, 0, comm);
    MV a = view(map, A, lda, p), v = view(map, V, ldv, q);
    Teuchos::SerialDenseMatrix<int, double> Bm(Teuchos::View, const_cast<double*>(B), ldb, p, q);
    MVT::MvTimesMatAddMv(alpha, a, Bm, beta, v);
    return 0;
  } catch (const std::exception& e) {
    return fail(e);
  }
}

extern "C" int ref_mv_dot(int64_t n, int k, const double* A, int64_t lda, const double* B,
                          int64_t ldb, double* out) {
  try {
    Epetra_SerialComm comm;
    Epetra_Map map((int)n, 0, comm);
    MV a = view(map, A, lda, k), b = view(map, B, ldb, k);
    std::vector<double> r(k);
    MVT::MvDot(a, b, r);
    std::memcpy(out, r.data(), sizeof(double) * k);
    return 0;
  } catch (const std::exception& e) {
    return fail(e);
  }
}

// type: 0 one, 1 two, 2 inf (Belos::NormType)
extern "C" int ref_mv_norm(int64_t n, int k, const double* A, int64_t lda, int type, double* out) {
  try {
    Epetra_SerialComm comm;
    Epetra_Map map((int)n, 0, comm);
    MV a = view(map, A, lda, k);
    std::vector<double> r(k);
    MVT::MvNorm(a, r, type == 0 ? Belos::OneNorm : (type == 1 ? Belos::TwoNorm : Belos::InfNorm));
    std::memcpy(out, r.data(), sizeof(double) * k);
    return 0;
  } catch (const std::exception& e) {
    return fail(e);
  }
}

extern "C" int ref_mv_add_mv(int64_t n, int k, double alpha, const double* A, int64_t lda,
                             double beta, const double* B, int64_t ldb, double* Y, int64_t ldy) {
  try {
    Epetra_SerialComm comm;
    Epetra_Map map((int)n, 0, comm);
    MV a = view(map, A, lda, k), b = view(map, B, ldb, k), y = view(map, Y, ldy, k);
    MVT::MvAddMv(alpha, a, beta, b, y);
    return 0;
  } catch (const std::exception& e) {
    return fail(e);
  }
}

extern "C" int ref_mv_scale_cols(int64_t n, int k, const double* alpha, double* Y, int64_t ldy) {
  try {
    Epetra_SerialComm comm;
    Epetra_Map map((int)n, 0, comm);
    MV y = view(map, Y, ldy, k);
    std::vector<double> a(alpha, alpha + k);
    MVT::MvScale(y, a);
    return 0;
  } catch (const std::exception& e) {
    return fail(e);
  }
}

// Epetra_MultiVector::Random with an explicit Epetra_Util seed; returns the advanced seed.
extern "C" unsigned int ref_mv_random(unsigned int seed, int64_t n, int k, double* Y, int64_t ldy) {
  Epetra_SerialComm comm;
  Epetra_Map map((int)n, 0, comm);
  MV y = view(map, Y, ldy, k);
  y.SetSeed(seed);
  MVT::MvRandom(y);
  return y.Seed();
}

// Full reference solve on a CSR matrix (serial Epetra): X is initial guess in / solution out.
extern "C" int ref_belos_solve_csr(const char* solver_name, int64_t n, const int32_t* rowptr,
                                   const int32_t* colind, const double* values, int numrhs,
                                   const double* Bh, int64_t ldb, double* Xh, int64_t ldx,
                                   const ref_gmres_params* p, ref_gmres_result* r);
struct ref_cheby_params;
static const ref_cheby_params* g_cheby = nullptr;  // set by ref_belos_solve_cheby_csr
// set by ref_belos_solve_hist_csr: native residual norms per block iteration ([row][rhs])
static double* g_hist = nullptr;
static int g_hist_cap = 0, g_hist_rows = 0;
namespace {
RCP<class Ifpack_Chebyshev> make_cheby(const RCP<Epetra_CrsMatrix>& A, const ref_cheby_params* c,
                                       double* lambda_used);
}

extern "C" int ref_block_gmres_csr(int64_t n, const int32_t* rowptr, const int32_t* colind,
                                   const double* values, int numrhs, const double* Bh, int64_t ldb,
                                   double* Xh, int64_t ldx, const ref_gmres_params* p,
                                   ref_gmres_result* r) {
  return ref_belos_solve_csr("BlockGmres", n, rowptr, colind, values, numrhs, Bh, ldb, Xh, ldx, p, r);
}

// the reference's solver managers on <double, Epetra_MultiVector, Epetra_Operator>, chosen by name
extern "C" int ref_belos_solve_csr(const char* solver_name, int64_t n, const int32_t* rowptr,
                                   const int32_t* colind, const double* values, int numrhs,
                                   const double* Bh, int64_t ldb, double* Xh, int64_t ldx,
                                   const ref_gmres_params* p, ref_gmres_result* r) {
  try {
    const std::string which(solver_name ? solver_name : "");
    std::memset(r, 0, sizeof(*r));
    Epetra_SerialComm comm;
    Epetra_Map map((int)n, 0, comm);
    RCP<Epetra_CrsMatrix> Acrs = csr_to_epetra(map, n, rowptr, colind, values);
    RCP<OP> A = Acrs;
    RCP<MV> B = rcp(new MV(View, map, const_cast<double*>(Bh), (int)ldb, numrhs));
    RCP<MV> X = rcp(new MV(View, map, Xh, (int)ldx, numrhs));
    int maxiters = p->max_iters;
    if (maxiters == -1) maxiters = (int)(n / std::max(1, p->block_size)) - 1;
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
    if (p->output_frequency > 0) belosList.set("Output Frequency", p->output_frequency);

    Belos::LinearProblem<ST, MV, OP> problem(A, X, B);
    RCP<Ifpack_Chebyshev> cheb;
    if (g_cheby) {  // right preconditioner, wrapped as the reference does (BelosEpetraAdapter.hpp:280-438)
      cheb = make_cheby(Acrs, g_cheby, nullptr);
      problem.setRightPrec(rcp(new Belos::EpetraPrecOp(cheb)));
    }
    if (!problem.setProblem()) {
      g_err = "Belos::LinearProblem failed to set up correctly";
      return -1;
    }
    Teuchos::TimeMonitor::zeroOutTimers();
    RCP<Belos::SolverManager<ST, MV, OP> > solver;
    RCP<Belos::B200ResidualHistory<ST, MV, OP> > rec;
    if (which == "BlockGmres") {
      RCP<Belos::BlockGmresSolMgr<ST, MV, OP> > bg =
          rcp(new Belos::BlockGmresSolMgr<ST, MV, OP>(rcp(&problem, false), rcp(&belosList, false)));
      if (g_hist_cap > 0) {  // setProblem() makes solve() rebuild the status-test combo with the recorder in it
        rec = rcp(new Belos::B200ResidualHistory<ST, MV, OP>());
        bg->setDebugStatusTest(rec);
        bg->setProblem(rcp(&problem, false));
      }
      solver = bg;
    } else if (which == "PseudoBlockGmres") {
      belosList.remove("Adaptive Block Size");
      belosList.remove("Flexible Gmres", false);
      solver = rcp(new Belos::PseudoBlockGmresSolMgr<ST, MV, OP>(rcp(&problem, false), rcp(&belosList, false)));
    } else if (which == "BlockCG") {
      belosList.remove("Num Blocks");
      belosList.remove("Maximum Restarts");
      belosList.remove("Flexible Gmres", false);
      solver = rcp(new Belos::BlockCGSolMgr<ST, MV, OP>(rcp(&problem, false), rcp(&belosList, false)));
    } else if (which == "PseudoBlockCG") {
      belosList.remove("Num Blocks");
      belosList.remove("Maximum Restarts");
      belosList.remove("Flexible Gmres", false);
      belosList.remove("Block Size");
      belosList.remove("Adaptive Block Size");
      belosList.remove("Orthogonalization");
      solver = rcp(new Belos::PseudoBlockCGSolMgr<ST, MV, OP>(rcp(&problem, false), rcp(&belosList, false)));
    } else {
      g_err = "unknown solver manager";
      return -1;
    }
    const auto t0 = std::chrono::steady_clock::now();
    const Belos::ReturnType ret = solver->solve();
    const auto t1 = std::chrono::steady_clock::now();
    r->converged = (ret == Belos::Converged) ? 1 : 0;
    r->num_iters = solver->getNumIters();
    r->achieved_tol = solver->achievedTol();
    r->loss_of_accuracy = solver->isLOADetected() ? 1 : 0;
    r->solve_seconds = std::chrono::duration<double>(t1 - t0).count();
    r->num_rhs = numrhs;
    if (nonnull(rec)) g_hist_rows = rec->copy_out(g_hist, g_hist_cap, numrhs);
    std::vector<double> actual(numrhs), rhs(numrhs);
    MV resid(map, numrhs);
    OPT::Apply(*A, *X, resid);
    MVT::MvAddMv(-1.0, resid, 1.0, *B, resid);
    MVT::MvNorm(resid, actual);
    MVT::MvNorm(*B, rhs);
    for (int i = 0; i < numrhs && i < 64; ++i)
      r->true_rel_resid[i] = rhs[i] > 0 ? actual[i] / rhs[i] : actual[i];
    return 0;
  } catch (const std::exception& e) {
    return fail(e);
  }
}

// BlockGmres with the native residual norms of every block iteration recorded (B200ResidualHistory
// attached through BlockGmresSolMgr::setDebugStatusTest): hist[row * numrhs + j], *rows = rows recorded
extern "C" int ref_belos_solve_hist_csr(int64_t n, const int32_t* rowptr, const int32_t* colind,
                                        const double* values, int numrhs, const double* Bh, int64_t ldb,
                                        double* Xh, int64_t ldx, const ref_gmres_params* p,
                                        ref_gmres_result* r, double* hist, int cap_rows, int* rows) {
  g_hist = hist;
  g_hist_cap = cap_rows;
  g_hist_rows = 0;
  const int rc = ref_belos_solve_csr("BlockGmres", n, rowptr, colind, values, numrhs, Bh, ldb, Xh, ldx, p, r);
  if (rows) *rows = g_hist_rows;
  g_hist = nullptr;
  g_hist_cap = 0;
  return rc;
}

// ---- Chebyshev right preconditioner: the reference's Ifpack_Chebyshev (ifpack/src/Ifpack_Chebyshev.cpp)
// in point mode, zero starting solution.  lambda_max < 0: estimated by the reference's own power
// method (Ifpack_Chebyshev::PowerMethod, :729-759) with an explicit RNG seed, then handed over as
// "chebyshev: max eigenvalue" so the run is reproducible.
struct ref_cheby_params {
  int degree;
  double lambda_max, eig_ratio, min_diag;
  int eig_iters;
  unsigned int seed;
};

namespace {
RCP<Ifpack_Chebyshev> make_cheby(const RCP<Epetra_CrsMatrix>& A, const ref_cheby_params* c,
                                 double* lambda_used) {
  double lam = c->lambda_max;
  if (lam < 0) {
    Epetra_Vector invd(A->RowMap());
    A->ExtractDiagonalCopy(invd);
    for (int i = 0; i < invd.MyLength(); ++i)
      invd[i] = (std::fabs(invd[i]) < c->min_diag) ? c->min_diag : 1.0 / invd[i];
    unsigned int seed = c->seed;
    if (Ifpack_Chebyshev::PowerMethod(*A, invd, c->eig_iters, lam, &seed) != 0)
      throw std::runtime_error("Ifpack_Chebyshev::PowerMethod failed");
  }
  Teuchos::ParameterList pl;
  pl.set("chebyshev: degree", c->degree);
  pl.set("chebyshev: max eigenvalue", lam);
  pl.set("chebyshev: ratio eigenvalue", c->eig_ratio);
  pl.set("chebyshev: min diagonal value", c->min_diag);
  pl.set("chebyshev: zero starting solution", true);
  RCP<Ifpack_Chebyshev> P = rcp(new Ifpack_Chebyshev(A.get()));
  if (P->SetParameters(pl) != 0 || P->Initialize() != 0 || P->Compute() != 0)
    throw std::runtime_error("Ifpack_Chebyshev set-up failed");
  if (lambda_used) *lambda_used = lam;
  return P;
}
}  // namespace

extern "C" int ref_cheby_apply(int64_t n, const int32_t* rowptr, const int32_t* colind,
                               const double* values, const ref_cheby_params* c, const double* X,
                               int64_t ldx, int k, double* Y, int64_t ldy, double* lambda_used) {
  try {
    Epetra_SerialComm comm;
    Epetra_Map map((int)n, 0, comm);
    RCP<Epetra_CrsMatrix> A = csr_to_epetra(map, n, rowptr, colind, values);
    RCP<Ifpack_Chebyshev> P = make_cheby(A, c, lambda_used);
    MV x = view(map, X, ldx, k), y = view(map, Y, ldy, k);
    return P->ApplyInverse(x, y);
  } catch (const std::exception& e) {
    return fail(e);
  }
}

extern "C" int ref_belos_solve_cheby_csr(const char* solver_name, int64_t n, const int32_t* rowptr,
                                         const int32_t* colind, const double* values,
                                         const ref_cheby_params* c, int numrhs, const double* Bh,
                                         int64_t ldb, double* Xh, int64_t ldx,
                                         const ref_gmres_params* p, ref_gmres_result* r) {
  g_cheby = c;
  const int rc = ref_belos_solve_csr(solver_name, n, rowptr, colind, values, numrhs, Bh, ldb, Xh, ldx, p, r);
  g_cheby = nullptr;
  return rc;
}

// ---- matrix ingest oracle: the reference's own readers.
// kind 0: Harwell-Boeing through Trilinos_Util_read_hb + the row insertion of
//         Belos::Util::createEpetraProblem (belos/epetra/src/BelosEpetraUtils.cpp:110-170, serial);
// kind 1: MatrixMarket through Trilinos_Util_ReadMatrixMarket2Epetra.
// Two-call protocol (rowptr == NULL: sizes only); the matrix is kept between the two calls.
namespace {
RCP<Epetra_CrsMatrix> g_read;
std::vector<double> g_read_b, g_read_x;
std::string g_read_key;
}  // namespace

extern "C" int ref_read_matrix(int kind, const char* path, int64_t* n, int64_t* nnz, int32_t* rowptr,
                               int32_t* colind, double* values, double* b, double* xexact) {
  try {
    static Epetra_SerialComm comm;
    const std::string key = std::to_string(kind) + ":" + path;
    if (key != g_read_key) {
      g_read = Teuchos::null;
      if (kind == 0) {
        int N = 0, nz = 0, *bindx = 0;
        double *val = 0, *xg = 0, *bb = 0, *xe = 0;
        Trilinos_Util_read_hb(const_cast<char*>(path), 0, &N, &nz, &val, &bindx, &xg, &bb, &xe);
        Epetra_Map map(N, 0, comm);
        std::vector<int> numnz(N);
        for (int i = 0; i < N; ++i) numnz[i] = bindx[i + 1] - bindx[i] + 1;
        g_read = rcp(new Epetra_CrsMatrix(Copy, map, numnz.data()));
        for (int i = 0; i < N; ++i) {
          g_read->InsertGlobalValues(i, bindx[i + 1] - bindx[i], val + bindx[i], bindx + bindx[i]);
          g_read->InsertGlobalValues(i, 1, val + i, &i);
        }
        g_read->FillComplete();
        g_read->OptimizeStorage();
        g_read_b.assign(bb, bb + N);
        g_read_x.assign(xe, xe + N);
        free(val); free(bindx); free(xg); free(bb); free(xe);
      } else {
        Epetra_Map* map = 0;
        Epetra_CrsMatrix* A = 0;
        Epetra_Vector *x = 0, *bv = 0, *xe = 0;
        if (Trilinos_Util_ReadMatrixMarket2Epetra(const_cast<char*>(path), comm, map, A, x, bv, xe) != 0)
          throw std::runtime_error("Trilinos_Util_ReadMatrixMarket2Epetra failed");
        A->OptimizeStorage();
        g_read = rcp(A);
        g_read_b.clear();
        g_read_x.clear();
        delete x; delete bv; delete xe;   // (the map must outlive the matrix: leaked on purpose)
      }
      g_read_key = key;
    }
    const int N = g_read->NumMyRows();
    *n = N;
    *nnz = g_read->NumMyNonzeros();
    if (!rowptr) return 0;
    int pos = 0;
    for (int i = 0; i < N; ++i) {
      int cnt = 0, *idx = 0;
      double* v = 0;
      g_read->ExtractMyRowView(i, cnt, v, idx);
      rowptr[i] = pos;
      for (int k = 0; k < cnt; ++k) {
        colind[pos] = g_read->GCID(idx[k]);
        values[pos++] = v[k];
      }
    }
    rowptr[N] = pos;
    if (b && !g_read_b.empty()) std::copy(g_read_b.begin(), g_read_b.end(), b);
    if (xexact && !g_read_x.empty()) std::copy(g_read_x.begin(), g_read_x.end(), xexact);
    return 0;
  } catch (const std::exception& e) {
    return fail(e);
  }
}

extern "C" int ref_timer_summary(char* buf, int buflen) {
  std::ostringstream os;
  Teuchos::TimeMonitor::summarize(os);
  const std::string s = os.str();
  std::strncpy(buf, s.c_str(), (size_t)buflen - 1);
  buf[buflen - 1] = 0;
  return 0;
}

// Thread control for the CPU baseline: Epetra's own OpenMP loops (EPETRA_HAVE_OMP) and the
// BLAS TPL (OpenBLAS).  n <= 1 gives the serial arithmetic order used for golden vectors.
#include <omp.h>
extern "C" void openblas_set_num_threads(int);
extern "C" int ref_set_threads(int n) {
  if (n < 1) n = 1;
  omp_set_num_threads(n);
  openblas_set_num_threads(n);
  return n;
}
extern "C" int ref_max_threads() { return omp_get_num_procs(); }
