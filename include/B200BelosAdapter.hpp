// B200BelosAdapter.hpp — the Belos plugin classes of the B200 path.
//
//   Belos::B200MultiVec : Belos::MultiVec<double>   (18 pure virtuals, belos/src/BelosMultiVec.hpp:107-210)
//   Belos::B200CrsOp    : Belos::Operator<double>   (belos/src/BelosOperator.hpp:118-145)
//
// An unmodified Belos::BlockGmresSolMgr<double, Belos::MultiVec<double>, Belos::Operator<double>>
// (with its DGKS / ICGS / IMGS managers and host-side Hessenberg least squares) drives these
// exactly as it drives Belos::EpetraMultiVec / Belos::EpetraOp
// (belos/epetra/src/BelosEpetraAdapter.hpp:116-261, .cpp:318-640).  Every virtual forwards to
// one C-ABI call of include/b200_belos.h; non-zero status becomes a Belos exception, as the
// reference adapter turns Epetra's non-zero return codes into EpetraMultiVecFailure /
// EpetraOpFailure (BelosEpetraAdapter.hpp:97-106).
#ifndef B200_BELOS_ADAPTER_HPP
#define B200_BELOS_ADAPTER_HPP

#include <iomanip>
#include <ostream>
#include <string>
#include <vector>

#include "BelosConfigDefs.hpp"
#include "BelosMultiVec.hpp"
#include "BelosOperator.hpp"
#include "BelosTypes.hpp"
#include "Teuchos_SerialDenseMatrix.hpp"
#include "b200_belos.h"

namespace Belos {

/// thrown on a non-zero status from a multivector entry point
class B200MultiVecFailure : public BelosError {
 public:
  B200MultiVecFailure(const std::string& what_arg) : BelosError(what_arg) {}
};
/// thrown on a non-zero status from the operator apply
class B200OpFailure : public BelosError {
 public:
  B200OpFailure(const std::string& what_arg) : BelosError(what_arg) {}
};

#define B200_MV_CHECK(call, who)                                                          \
  do {                                                                                    \
    const int b200_rc_ = (call);                                                          \
    TEUCHOS_TEST_FOR_EXCEPTION(b200_rc_ != 0, B200MultiVecFailure,                        \
                               "Belos::B200MultiVec::" who ": " #call " returned "        \
                                   << b200_rc_ << ": " << b200_last_error());             \
  } while (0)

class B200MultiVec : public MultiVec<double> {
 public:
  /// n_local rows of an n_global-row multivector starting at global row row_offset
  /// (Epetra_MultiVector(Map, numvecs, zeroOut), epetra/src/Epetra_MultiVector.cpp:60-80)
  B200MultiVec(b200_ctx* ctx, ptrdiff_t n_local, ptrdiff_t n_global, ptrdiff_t row_offset,
               int numvecs, bool zeroOut = true)
      : h_(nullptr) {
    B200_MV_CHECK(b200_mv_create(ctx, n_local, n_global, row_offset, numvecs, zeroOut ? 1 : 0, &h_),
                  "B200MultiVec");
  }
  /// adopt an existing handle (takes ownership)
  explicit B200MultiVec(b200_mv* h) : h_(h) {}
  B200MultiVec(const B200MultiVec&) = delete;
  B200MultiVec& operator=(const B200MultiVec&) = delete;
  ~B200MultiVec() override { b200_mv_destroy(h_); }

  b200_mv* handle() const { return h_; }

  MultiVec<double>* Clone(const int numvecs) const override {
    b200_mv* c = nullptr;
    B200_MV_CHECK(b200_mv_clone(h_, numvecs, &c), "Clone");
    return new B200MultiVec(c);
  }
  MultiVec<double>* CloneCopy() const override {
    b200_mv* c = nullptr;
    B200_MV_CHECK(b200_mv_clone_copy(h_, nullptr, 0, &c), "CloneCopy");
    return new B200MultiVec(c);
  }
  MultiVec<double>* CloneCopy(const std::vector<int>& index) const override {
    b200_mv* c = nullptr;
    B200_MV_CHECK(b200_mv_clone_copy(h_, index.data(), (int)index.size(), &c), "CloneCopy");
    return new B200MultiVec(c);
  }
  MultiVec<double>* CloneViewNonConst(const std::vector<int>& index) override {
    b200_mv* c = nullptr;
    B200_MV_CHECK(b200_mv_clone_view(h_, index.data(), (int)index.size(), &c), "CloneViewNonConst");
    return new B200MultiVec(c);
  }
  const MultiVec<double>* CloneView(const std::vector<int>& index) const override {
    b200_mv* c = nullptr;
    B200_MV_CHECK(b200_mv_clone_view(h_, index.data(), (int)index.size(), &c), "CloneView");
    return new B200MultiVec(c);
  }

  ptrdiff_t GetGlobalLength() const override { return (ptrdiff_t)b200_mv_global_length(h_); }
  int GetNumberVecs() const override { return b200_mv_num_vecs(h_); }

  void MvTimesMatAddMv(const double alpha, const MultiVec<double>& A,
                       const Teuchos::SerialDenseMatrix<int, double>& B,
                       const double beta) override {
    const B200MultiVec* a = cast(A, "MvTimesMatAddMv");
    B200_MV_CHECK(b200_mv_times_mat_add_mv(h_, alpha, a->h_, B.values(), B.stride(), B.numRows(),
                                           B.numCols(), beta),
                  "MvTimesMatAddMv");
  }
  void MvAddMv(const double alpha, const MultiVec<double>& A, const double beta,
               const MultiVec<double>& B) override {
    const B200MultiVec* a = cast(A, "MvAddMv");
    const B200MultiVec* b = cast(B, "MvAddMv");
    B200_MV_CHECK(b200_mv_add_mv(h_, alpha, a->h_, beta, b->h_), "MvAddMv");
  }
  void MvScale(const double alpha) override { B200_MV_CHECK(b200_mv_scale(h_, alpha), "MvScale"); }
  void MvScale(const std::vector<double>& alpha) override {
    B200_MV_CHECK(b200_mv_scale_cols(h_, alpha.data(), (int)alpha.size()), "MvScale");
  }
  void MvTransMv(const double alpha, const MultiVec<double>& A,
                 Teuchos::SerialDenseMatrix<int, double>& B) const override {
    const B200MultiVec* a = cast(A, "MvTransMv");
    B200_MV_CHECK(b200_mv_trans_mv(h_, alpha, a->h_, B.values(), B.stride(), B.numRows(),
                                   B.numCols()),
                  "MvTransMv");
  }
  void MvDot(const MultiVec<double>& A, std::vector<double>& b) const override {
    const B200MultiVec* a = cast(A, "MvDot");
    B200_MV_CHECK(b200_mv_dot(h_, a->h_, b.data(), (int)b.size()), "MvDot");
  }
  void MvNorm(std::vector<double>& normvec, NormType type = TwoNorm) const override {
    int t;
    switch (type) {
      case OneNorm: t = B200_NORM_ONE; break;
      case TwoNorm: t = B200_NORM_TWO; break;
      case InfNorm: t = B200_NORM_INF; break;
      default:
        TEUCHOS_TEST_FOR_EXCEPTION(true, std::invalid_argument,
                                   "Belos::B200MultiVec::MvNorm: Invalid norm_type "
                                       << type << ".  The current list of valid norm types is "
                                       "{OneNorm, TwoNorm, InfNorm}.");
    }
    B200_MV_CHECK(b200_mv_norm(h_, t, normvec.data(), (int)normvec.size()), "MvNorm");
  }
  void SetBlock(const MultiVec<double>& A, const std::vector<int>& index) override {
    const B200MultiVec* a = cast(A, "SetBlock");
    B200_MV_CHECK(b200_mv_set_block(h_, a->h_, index.data(), (int)index.size()), "SetBlock");
  }
  void MvRandom() override { B200_MV_CHECK(b200_mv_random(h_), "MvRandom"); }
  void MvInit(const double alpha) override { B200_MV_CHECK(b200_mv_init(h_, alpha), "MvInit"); }
  void MvPrint(std::ostream& os) const override {
    const ptrdiff_t n = (ptrdiff_t)b200_mv_local_length(h_);
    const int k = GetNumberVecs();
    std::vector<double> host((size_t)(n > 0 ? n : 1) * (k > 0 ? k : 1));
    if (k > 0) B200_MV_CHECK(b200_mv_copy_to_host(h_, host.data(), n > 0 ? n : 1), "MvPrint");
    os << "B200MultiVec: global length " << GetGlobalLength() << ", local length " << n << ", "
       << k << " vectors" << std::endl;
    for (ptrdiff_t i = 0; i < n; ++i) {
      for (int j = 0; j < k; ++j) os << std::setw(22) << std::setprecision(15) << host[(size_t)j * n + i];
      os << std::endl;
    }
  }

#ifdef HAVE_BELOS_TSQR
  /// TSQR hooks (belos/src/BelosMultiVec.hpp:213-293): Cholesky QR of the block on the Gram / update kernels,
  /// SVD of the small factor on the host (trilinos_b200/csrc/tsqr.cu)
  void factorExplicit(MultiVec<double>& Q, Teuchos::SerialDenseMatrix<int, double>& R,
                      const bool forceNonnegativeDiagonal = false) override {
    B200MultiVec* q = dynamic_cast<B200MultiVec*>(&Q);
    TEUCHOS_TEST_FOR_EXCEPTION(q == nullptr, B200MultiVecFailure,
                               "Belos::B200MultiVec::factorExplicit: Q is not a Belos::B200MultiVec.");
    const int b = GetNumberVecs();
    TEUCHOS_TEST_FOR_EXCEPTION(R.numRows() < b || R.numCols() < b, std::invalid_argument,
                               "Belos::B200MultiVec::factorExplicit: R is " << R.numRows() << " x " << R.numCols()
                                                                             << " but the block has " << b << " columns.");
    B200_MV_CHECK(b200_mv_factor_explicit(h_, q->h_, R.values(), R.stride(), forceNonnegativeDiagonal ? 1 : 0),
                  "factorExplicit");
  }
  int revealRank(Teuchos::SerialDenseMatrix<int, double>& R, const double& tol) override {
    int rank = 0;
    B200_MV_CHECK(b200_mv_reveal_rank(h_, R.values(), R.stride(), tol, &rank), "revealRank");
    return rank;
  }
#endif  // HAVE_BELOS_TSQR

  /// host transfer helpers for drivers (ExtractCopy / copy-in)
  void copyToHost(double* host, ptrdiff_t ldh) const {
    B200_MV_CHECK(b200_mv_copy_to_host(h_, host, ldh), "copyToHost");
  }
  void copyFromHost(const double* host, ptrdiff_t ldh) {
    B200_MV_CHECK(b200_mv_copy_from_host(h_, host, ldh), "copyFromHost");
  }

 private:
  static const B200MultiVec* cast(const MultiVec<double>& A, const char* who) {
    const B200MultiVec* a = dynamic_cast<const B200MultiVec*>(&A);
    TEUCHOS_TEST_FOR_EXCEPTION(a == nullptr, B200MultiVecFailure,
                               "Belos::B200MultiVec::" << who
                                                       << ": cast from Belos::MultiVec<> to "
                                                          "Belos::B200MultiVec failed.");
    return a;
  }
  b200_mv* h_;
};

class B200CrsOp : public Operator<double> {
 public:
  /// does not take ownership of the operator handle unless own == true
  explicit B200CrsOp(b200_op* op, bool own = false) : op_(op), own_(own) {}
  ~B200CrsOp() override {
    if (own_) b200_op_destroy(op_);
  }
  void Apply(const MultiVec<double>& x, MultiVec<double>& y, ETrans trans = NOTRANS) const override {
    const B200MultiVec* vx = dynamic_cast<const B200MultiVec*>(&x);
    B200MultiVec* vy = dynamic_cast<B200MultiVec*>(&y);
    TEUCHOS_TEST_FOR_EXCEPTION(vx == nullptr || vy == nullptr, B200OpFailure,
                               "Belos::B200CrsOp::Apply: x and/or y could not be dynamic cast to "
                               "a Belos::B200MultiVec.");
    const int info = b200_op_apply(op_, vx->handle(), vy->handle(), trans == NOTRANS ? 0 : 1);
    TEUCHOS_TEST_FOR_EXCEPTION(info != 0, B200OpFailure,
                               "Belos::B200CrsOp::Apply: b200_op_apply returned a nonzero error "
                               "code of " << info << ": " << b200_last_error());
  }
  bool HasApplyTranspose() const override { return false; }

 private:
  b200_op* op_;
  bool own_;
};

/// Chebyshev smoother (the reference's Ifpack_Chebyshev, point mode) as a Belos::Operator whose
/// Apply is ApplyInverse -- the role Belos::EpetraPrecOp plays for Epetra preconditioners
/// (belos/epetra/src/BelosEpetraAdapter.hpp:280-438), to be handed to
/// LinearProblem::setRightPrec / setLeftPrec.
class B200ChebyshevOp : public Operator<double> {
 public:
  explicit B200ChebyshevOp(b200_cheby* c, bool own = false) : c_(c), own_(own) {}
  ~B200ChebyshevOp() override {
    if (own_) b200_cheby_destroy(c_);
  }
  void Apply(const MultiVec<double>& x, MultiVec<double>& y, ETrans trans = NOTRANS) const override {
    const B200MultiVec* vx = dynamic_cast<const B200MultiVec*>(&x);
    B200MultiVec* vy = dynamic_cast<B200MultiVec*>(&y);
    TEUCHOS_TEST_FOR_EXCEPTION(vx == nullptr || vy == nullptr, B200OpFailure,
                               "Belos::B200ChebyshevOp::Apply: x and/or y could not be dynamic "
                               "cast to a Belos::B200MultiVec.");
    TEUCHOS_TEST_FOR_EXCEPTION(trans != NOTRANS, B200OpFailure,
                               "Belos::B200ChebyshevOp::Apply: transpose is not available.");
    const int info = b200_cheby_apply(c_, vx->handle(), vy->handle());
    TEUCHOS_TEST_FOR_EXCEPTION(info != 0, B200OpFailure,
                               "Belos::B200ChebyshevOp::Apply: b200_cheby_apply returned a nonzero "
                               "error code of " << info << ": " << b200_last_error());
  }
  bool HasApplyTranspose() const override { return false; }

 private:
  b200_cheby* c_;
  bool own_;
};

}  // namespace Belos

#endif
