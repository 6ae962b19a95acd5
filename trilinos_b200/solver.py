"""Driver for the unmodified Belos::BlockGmresSolMgr on the B200 plugin classes
(libb200belos_solver.so, include/b200_belos_solver.h).

Parameter names mirror the Belos ParameterList keys of
belos/src/BelosBlockGmresSolMgr.hpp:430-490 as the reference driver sets them
(belos/epetra/example/BlockGmres/BlockGmresEpetraGaleriEx.cpp:170-190).
"""
import ctypes as C
import os

from ._lib import B200Error, SOLVER_LIB_PATH, check, lib, vp

# Belos::MsgType (belos/src/BelosTypes.hpp)
ERRORS, WARNINGS, ITERATION_DETAILS, ORTHO_DETAILS, FINAL_SUMMARY, TIMING_DETAILS, STATUS_TEST_DETAILS, DEBUG = (
    0, 0x1, 0x2, 0x4, 0x8, 0x10, 0x20, 0x40)


class GmresParams(C.Structure):
    _fields_ = [("block_size", C.c_int), ("num_blocks", C.c_int), ("max_iters", C.c_int),
                ("max_restarts", C.c_int), ("tol", C.c_double), ("ortho", C.c_char_p),
                ("verbosity", C.c_int), ("output_frequency", C.c_int), ("adaptive_block_size", C.c_int),
                ("flexible", C.c_int), ("show_timers", C.c_int)]


class GmresResult(C.Structure):
    _fields_ = [("converged", C.c_int), ("num_iters", C.c_int), ("achieved_tol", C.c_double),
                ("loss_of_accuracy", C.c_int), ("solve_seconds", C.c_double), ("num_rhs", C.c_int),
                ("true_rel_resid", C.c_double * 64)]


_slib = None


def solver_lib():
    """Loads libb200belos_solver.so (fails loudly if it was not built)."""
    global _slib
    if _slib is None:
        if not os.path.exists(SOLVER_LIB_PATH):
            raise ImportError(f"{SOLVER_LIB_PATH} not found: run __graft_entry__.build() where the reference "
                              "tree is available (the Belos headers are compiled from it)")
        os.environ.setdefault("OPENBLAS_NUM_THREADS", "1")  # host side is k x k dense algebra only
        _slib = C.CDLL(SOLVER_LIB_PATH, mode=C.RTLD_GLOBAL)
        _slib.b200_gmres_default_params.argtypes = [C.POINTER(GmresParams)]
        _slib.b200_belos_block_gmres_solve.argtypes = [vp, vp, vp, C.POINTER(GmresParams), C.POINTER(GmresResult)]
        _slib.b200_belos_block_gmres_solve.restype = C.c_int
        _slib.b200_belos_solve.argtypes = [C.c_char_p, vp, vp, vp, C.POINTER(GmresParams), C.POINTER(GmresResult)]
        _slib.b200_belos_solve.restype = C.c_int
        _slib.b200_belos_solve_prec.argtypes = [C.c_char_p, vp, vp, C.c_int, vp, vp, C.POINTER(GmresParams),
                                                C.POINTER(GmresResult)]
        _slib.b200_belos_solve_prec.restype = C.c_int
        _slib.b200_belos_solve_hist.argtypes = [C.c_char_p, vp, vp, C.c_int, vp, vp, C.POINTER(GmresParams),
                                                C.POINTER(GmresResult), C.POINTER(C.c_double), C.c_int,
                                                C.POINTER(C.c_int)]
        _slib.b200_belos_solve_hist.restype = C.c_int
        _slib.b200_belos_mvop_test.argtypes = [vp, C.c_int, C.c_int, C.c_int]
        _slib.b200_belos_mvop_test.restype = C.c_int
        _slib.b200_belos_ortho_test.argtypes = [vp, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                                C.POINTER(C.c_int)]
        _slib.b200_belos_ortho_test.restype = C.c_int
        _slib.b200_belos_timer_summary.argtypes = [C.c_char_p, C.c_int]
    return _slib


SOLVER_EXPORTED = ("b200_gmres_default_params", "b200_belos_block_gmres_solve", "b200_belos_solve", "b200_belos_solve_prec",
                   "b200_belos_solve_hist", "b200_belos_mvop_test", "b200_belos_ortho_test", "b200_belos_timer_summary")


class BlockGmresSolMgr:
    """Belos::BlockGmresSolMgr<double, Belos::MultiVec<double>, Belos::Operator<double>>."""
    SOLVER = b"BlockGmres"

    def __init__(self, op, block_size=1, num_blocks=300, max_iters=1000, max_restarts=20, tol=1e-8,
                 ortho="DGKS", verbosity=0, output_frequency=-1, adaptive_block_size=True, flexible=False,
                 show_timers=False, right_prec=None, left_prec=None):
        self.op = op
        # LinearProblem::setRightPrec / setLeftPrec with a trilinos_b200.Chebyshev
        self.prec, self.left = (left_prec, 1) if left_prec is not None else (right_prec, 0)
        self._ortho = ortho.encode()
        self.params = GmresParams(block_size, num_blocks, max_iters, max_restarts, tol, self._ortho, verbosity,
                                  output_frequency, 1 if adaptive_block_size else 0, 1 if flexible else 0,
                                  1 if show_timers else 0)
        self.result = None

    def solve(self, X, B, history=0):
        """X: initial guess in / solution out; B: right-hand sides.  Returns a dict.  history > 0 (BlockGmres only):
        additionally record the native residual norms before each of the first `history` block iterations
        (b200_belos_solve_hist) under the key "native_resid_history" ([iteration][rhs])."""
        r = GmresResult()
        prec = self.prec._h if self.prec is not None else None
        hist = None
        if history > 0:
            k = B.GetNumberVecs()
            hist = (C.c_double * (history * k))()
            rows = C.c_int(0)
            rc = solver_lib().b200_belos_solve_hist(self.SOLVER, self.op._h, prec, self.left, X._h, B._h,
                                                    C.byref(self.params), C.byref(r), hist, history, C.byref(rows))
        else:
            rc = solver_lib().b200_belos_solve_prec(self.SOLVER, self.op._h, prec, self.left, X._h, B._h,
                                                    C.byref(self.params), C.byref(r))
        check(rc)
        self.result = dict(converged=bool(r.converged), num_iters=r.num_iters, achieved_tol=r.achieved_tol,
                           loss_of_accuracy=bool(r.loss_of_accuracy), solve_seconds=r.solve_seconds,
                           true_rel_resid=[r.true_rel_resid[i] for i in range(min(r.num_rhs, 64))])
        if hist is not None:
            nrow = min(rows.value, history)
            self.result["native_resid_history"] = [[hist[i * k + j] for j in range(k)] for i in range(nrow)]
        return self.result

    def getNumIters(self):
        return self.result["num_iters"]

    def achievedTol(self):
        return self.result["achieved_tol"]

    def isLOADetected(self):
        return self.result["loss_of_accuracy"]

    @staticmethod
    def timer_summary():
        buf = C.create_string_buffer(1 << 16)
        solver_lib().b200_belos_timer_summary(buf, len(buf))
        return buf.value.decode(errors="replace")


class PseudoBlockGmresSolMgr(BlockGmresSolMgr):
    """Belos::PseudoBlockGmresSolMgr (belos/src/BelosPseudoBlockGmresSolMgr.hpp): block_size independent
    single-vector GMRES recurrences advanced together; the same plugin kernels."""
    SOLVER = b"PseudoBlockGmres"


class BlockCGSolMgr(BlockGmresSolMgr):
    """Belos::BlockCGSolMgr (belos/src/BelosBlockCGSolMgr.hpp); num_blocks / max_restarts are ignored."""
    SOLVER = b"BlockCG"


class PseudoBlockCGSolMgr(BlockGmresSolMgr):
    """Belos::PseudoBlockCGSolMgr (belos/src/BelosPseudoBlockCGSolMgr.hpp)."""
    SOLVER = b"PseudoBlockCG"


SOLVER_MANAGERS = {"BlockGmres": BlockGmresSolMgr, "PseudoBlockGmres": PseudoBlockGmresSolMgr,
                   "BlockCG": BlockCGSolMgr, "PseudoBlockCG": PseudoBlockCGSolMgr}


def mvop_test(ctx, n=100, ncols=5, verbose=False):
    """Belos::TestMultiVecTraits + TestOperatorTraits (belos/src/BelosMVOPTester.hpp) on the B200 classes."""
    rc = solver_lib().b200_belos_mvop_test(ctx._h, n, ncols, 1 if verbose else 0)
    if rc != 0:
        raise B200Error(rc, lib.b200_last_error().decode(errors="replace"))
    return True


def ortho_test(ctx, ortho="TSQR", n=500, ncols=12, size_x1=3, size_x2=5, rank_revealing=True, verbose=False):
    """Belos::Test::OrthoManagerTester::runTests (belos/src/BelosOrthoManagerTest.hpp) on the B200 classes with the
    orthogonalisation manager named `ortho`; returns the number of failed checks."""
    nf = C.c_int(-1)
    rc = solver_lib().b200_belos_ortho_test(ctx._h, ortho.encode(), n, ncols, size_x1, size_x2, 1 if rank_revealing else 0,
                                            1 if verbose else 0, C.byref(nf))
    if rc != 0:
        raise B200Error(rc, lib.b200_last_error().decode(errors="replace"))
    return nf.value
