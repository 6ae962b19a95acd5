// Belos::StatusTest that records the native residual norms of every block iteration.
//
// Attached through BlockGmresSolMgr::setDebugStatusTest (belos/src/BelosBlockGmresSolMgr.hpp:240,
// :883-886: the test joins the manager's OR combo, which evaluates every member each iteration,
// BelosStatusTestCombo.hpp orOp).  It never passes, so it cannot change the iteration; it only
// copies what Iteration::getNativeResiduals returns (for BlockGmresIter: the 2-norms of the
// trailing b entries of the rotated right-hand side z, BelosBlockGmresIter.hpp:542-557 -- host data,
// no extra kernel).  Templated on the multivector / operator types so that the same class records
// the reference's Epetra run (oracle/ref_harness.cpp) and the B200 run (belos_solver.cpp); the two
// histories are what the full-size parity tests compare.
#ifndef B200_RESIDUAL_HISTORY_HPP
#define B200_RESIDUAL_HISTORY_HPP

#include <vector>

#include "BelosIteration.hpp"
#include "BelosStatusTest.hpp"

namespace Belos {

template <class ScalarType, class MV, class OP>
class B200ResidualHistory : public StatusTest<ScalarType, MV, OP> {
 public:
  typedef typename Teuchos::ScalarTraits<ScalarType>::magnitudeType MagnitudeType;

  StatusType checkStatus(Iteration<ScalarType, MV, OP>* iSolver) override {
    const int it = iSolver->getNumIters();
    if (it == last_iter_) return Failed;  // the manager re-checks after a restart: one row per iteration
    last_iter_ = it;
    std::vector<MagnitudeType> norms;
    iSolver->getNativeResiduals(&norms);
    iters_.push_back(it);
    width_.push_back((int)norms.size());
    for (MagnitudeType v : norms) hist_.push_back(v);
    return Failed;
  }
  StatusType getStatus() const override { return Failed; }
  void reset() override {}
  void print(std::ostream& os, int indent = 0) const override {
    for (int j = 0; j < indent; ++j) os << ' ';
    os << "B200ResidualHistory: " << iters_.size() << " rows recorded" << std::endl;
  }

  // row r: iteration number iters()[r], norms values()[offset .. offset + widths()[r])
  const std::vector<int>& iters() const { return iters_; }
  const std::vector<int>& widths() const { return width_; }
  const std::vector<MagnitudeType>& values() const { return hist_; }

  // flat copy: out[r * ncols + j] for the first min(rows, cap_rows) rows; returns the row count
  int copy_out(double* out, int cap_rows, int ncols) const {
    const int rows = (int)iters_.size();
    size_t off = 0;
    for (int r = 0; r < rows; ++r) {
      if (out && r < cap_rows)
        for (int j = 0; j < ncols; ++j) out[(size_t)r * ncols + j] = j < width_[r] ? (double)hist_[off + j] : 0.0;
      off += (size_t)width_[r];
    }
    return rows;
  }

 private:
  int last_iter_ = -1;
  std::vector<int> iters_, width_;
  std::vector<MagnitudeType> hist_;
};

}  // namespace Belos

#endif
