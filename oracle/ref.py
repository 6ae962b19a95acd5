"""ctypes wrapper of oracle/_ref/libref_epetra_belos.so — the REFERENCE's own compiled Epetra /
Galeri / Belos code (built by oracle/Makefile.ref from the sources under /root/reference).

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): golden-vector generation, parity checks and
the CPU baseline of bench.py.  `available()` is False where the library was not built.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(_HERE, "_ref", "libref_epetra_belos.so")
_lib = None


class GmresParams(C.Structure):
    _fields_ = [("block_size", C.c_int), ("num_blocks", C.c_int), ("max_iters", C.c_int),
                ("max_restarts", C.c_int), ("tol", C.c_double), ("ortho", C.c_char_p),
                ("verbosity", C.c_int), ("output_frequency", C.c_int), ("adaptive_block_size", C.c_int),
                ("flexible", C.c_int), ("show_timers", C.c_int)]


class GmresResult(C.Structure):
    _fields_ = [("converged", C.c_int), ("num_iters", C.c_int), ("achieved_tol", C.c_double),
                ("loss_of_accuracy", C.c_int), ("solve_seconds", C.c_double), ("num_rhs", C.c_int),
                ("true_rel_resid", C.c_double * 64)]


class ChebyParams(C.Structure):
    """ifpack/src/Ifpack_Chebyshev.cpp:160-175 parameter names: degree, max eigenvalue (-1: power method), ratio
    eigenvalue, min diagonal value, eigenvalue max iterations; seed pins the power method's start vector."""
    _fields_ = [("degree", C.c_int), ("lambda_max", C.c_double), ("eig_ratio", C.c_double), ("min_diag", C.c_double),
                ("eig_iters", C.c_int), ("seed", C.c_uint)]


def available():
    return os.path.exists(SO)


def lib():
    global _lib
    if _lib is None:
        os.environ.setdefault("OPENBLAS_NUM_THREADS", "1")
        os.environ.setdefault("OMP_NUM_THREADS", "1")
        _lib = C.CDLL(SO, mode=os.RTLD_LOCAL | os.RTLD_NOW)
        _lib.ref_last_error.restype = C.c_char_p
        _lib.ref_mv_random.restype = C.c_uint
        _lib.ref_set_threads(1)
    return _lib


def set_threads(n):
    return lib().ref_set_threads(int(n))


def max_threads():
    return lib().ref_max_threads()


def _ck(rc):
    if rc != 0:
        raise RuntimeError("reference: " + lib().ref_last_error().decode(errors="replace"))


def _d(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _i32(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


def _F(a):
    a = np.asarray(a, dtype=np.float64)
    if a.ndim == 1:
        a = a.reshape(-1, 1)
    return np.asfortranarray(a)


def galeri(name, nx, ny=1, nz=1, conv=1.0, diff=1.0e-5):
    """Galeri::CreateCrsMatrix(name) on Galeri::CreateMap("Linear"), serial; CSR with global ids."""
    nnz = C.c_int64()
    nm = name.encode()
    _ck(lib().ref_galeri(nm, nx, ny, nz, C.c_double(conv), C.c_double(diff), C.byref(nnz), None, None, None))
    n = nx * ny * nz
    rp = np.zeros(n + 1, dtype=np.int32)
    ci = np.zeros(nnz.value, dtype=np.int32)
    va = np.zeros(nnz.value)
    _ck(lib().ref_galeri(nm, nx, ny, nz, C.c_double(conv), C.c_double(diff), C.byref(nnz), _i32(rp), _i32(ci), _d(va)))
    return n, rp, ci.astype(np.int64), va


def _csr(rowptr, colind, values):
    return (np.ascontiguousarray(rowptr, dtype=np.int32), np.ascontiguousarray(colind, dtype=np.int32),
            np.ascontiguousarray(values, dtype=np.float64))


def apply(rowptr, colind, values, X):
    rp, ci, va = _csr(rowptr, colind, values)
    X = _F(X)
    n = len(rp) - 1
    Y = np.zeros((n, X.shape[1]), order="F")
    _ck(lib().ref_apply(C.c_int64(n), _i32(rp), _i32(ci), _d(va), _d(X), C.c_int64(n), X.shape[1], _d(Y), C.c_int64(n)))
    return Y


def trans_mv(alpha, A, V):
    A, V = _F(A), _F(V)
    n, p, q = A.shape[0], A.shape[1], V.shape[1]
    Cm = np.zeros((p, q), order="F")
    _ck(lib().ref_mv_trans_mv(C.c_int64(n), p, q, C.c_double(alpha), _d(A), C.c_int64(n), _d(V), C.c_int64(n), _d(Cm), p))
    return Cm


def times_mat_add_mv(alpha, A, B, beta, V):
    A, B = _F(A), _F(B)
    V = _F(V).copy(order="F")
    n, p, q = A.shape[0], A.shape[1], V.shape[1]
    _ck(lib().ref_mv_times_mat_add_mv(C.c_int64(n), p, q, C.c_double(alpha), _d(A), C.c_int64(n), _d(B), p,
                                      C.c_double(beta), _d(V), C.c_int64(n)))
    return V


def dot(A, B):
    A, B = _F(A), _F(B)
    out = np.zeros(A.shape[1])
    _ck(lib().ref_mv_dot(C.c_int64(A.shape[0]), A.shape[1], _d(A), C.c_int64(A.shape[0]), _d(B), C.c_int64(B.shape[0]), _d(out)))
    return out


def norm(A, norm_type=1):
    A = _F(A)
    out = np.zeros(A.shape[1])
    _ck(lib().ref_mv_norm(C.c_int64(A.shape[0]), A.shape[1], _d(A), C.c_int64(A.shape[0]), norm_type, _d(out)))
    return out


def add_mv(alpha, A, beta, B):
    A, B = _F(A), _F(B)
    Y = np.zeros_like(A, order="F")
    n = A.shape[0]
    _ck(lib().ref_mv_add_mv(C.c_int64(n), A.shape[1], C.c_double(alpha), _d(A), C.c_int64(n), C.c_double(beta), _d(B),
                            C.c_int64(n), _d(Y), C.c_int64(n)))
    return Y


def scale_cols(alpha, Y):
    Y = _F(Y).copy(order="F")
    a = np.ascontiguousarray(np.broadcast_to(np.asarray(alpha, dtype=np.float64), (Y.shape[1],)))
    _ck(lib().ref_mv_scale_cols(C.c_int64(Y.shape[0]), Y.shape[1], _d(a), _d(Y), C.c_int64(Y.shape[0])))
    return Y


def random(seed, n, k):
    Y = np.zeros((n, k), order="F")
    s = lib().ref_mv_random(C.c_uint(seed), C.c_int64(n), k, _d(Y), C.c_int64(n))
    return Y, int(s)


def block_gmres(rowptr, colind, values, B, X0=None, block_size=1, num_blocks=300, max_iters=-1, max_restarts=20,
                tol=1e-8, ortho="DGKS", adaptive_block_size=True, show_timers=False, verbosity=0, solver="BlockGmres"):
    """Belos::BlockGmresSolMgr<double,Epetra_MultiVector,Epetra_Operator> on serial Epetra."""
    rp, ci, va = _csr(rowptr, colind, values)
    B = _F(B)
    n, k = B.shape
    X = np.zeros((n, k), order="F") if X0 is None else _F(X0).copy(order="F")
    o = ortho.encode()
    p = GmresParams(block_size, num_blocks, max_iters, max_restarts, tol, o, verbosity, -1,
                    1 if adaptive_block_size else 0, 0, 1 if show_timers else 0)
    r = GmresResult()
    _ck(lib().ref_belos_solve_csr(solver.encode(), C.c_int64(n), _i32(rp), _i32(ci), _d(va), k, _d(B), C.c_int64(n), _d(X), C.c_int64(n),
                                  C.byref(p), C.byref(r)))
    return X, dict(converged=bool(r.converged), num_iters=r.num_iters, achieved_tol=r.achieved_tol,
                   loss_of_accuracy=bool(r.loss_of_accuracy), solve_seconds=r.solve_seconds,
                   true_rel_resid=[r.true_rel_resid[i] for i in range(min(k, 64))])


def block_gmres_history(rowptr, colind, values, B, X0=None, block_size=1, num_blocks=300, max_iters=-1, max_restarts=20,
                        tol=1e-8, ortho="DGKS", max_rows=100000):
    """block_gmres() with the native residual norms of every block iteration recorded (Belos::B200ResidualHistory
    attached through BlockGmresSolMgr::setDebugStatusTest): returns (X, result dict, hist[rows, k])."""
    rp, ci, va = _csr(rowptr, colind, values)
    B = _F(B)
    n, k = B.shape
    X = np.zeros((n, k), order="F") if X0 is None else _F(X0).copy(order="F")
    o = ortho.encode()
    p = GmresParams(block_size, num_blocks, max_iters, max_restarts, tol, o, 0, -1, 1, 0, 0)
    r = GmresResult()
    hist = np.zeros((max_rows, k))
    rows = C.c_int(0)
    _ck(lib().ref_belos_solve_hist_csr(C.c_int64(n), _i32(rp), _i32(ci), _d(va), k, _d(B), C.c_int64(n), _d(X), C.c_int64(n),
                                       C.byref(p), C.byref(r), _d(hist), max_rows, C.byref(rows)))
    return X, dict(converged=bool(r.converged), num_iters=r.num_iters, achieved_tol=r.achieved_tol,
                   loss_of_accuracy=bool(r.loss_of_accuracy), solve_seconds=r.solve_seconds,
                   true_rel_resid=[r.true_rel_resid[i] for i in range(min(k, 64))]), hist[:min(rows.value, max_rows)].copy()


def cheby_apply(rowptr, colind, values, X, degree=3, lambda_max=-1.0, eig_ratio=30.0, min_diag=0.0, eig_iters=10, seed=1):
    """Y = Ifpack_Chebyshev(A).ApplyInverse(X) (point mode, zero starting solution); returns (Y, lambda_max used)."""
    rp, ci, va = _csr(rowptr, colind, values)
    X = _F(X)
    n, k = X.shape
    Y = np.zeros((n, k), order="F")
    cp = ChebyParams(degree, lambda_max, eig_ratio, min_diag, eig_iters, seed)
    lam = C.c_double(0.0)
    _ck(lib().ref_cheby_apply(C.c_int64(n), _i32(rp), _i32(ci), _d(va), C.byref(cp), _d(X), C.c_int64(n), k, _d(Y),
                              C.c_int64(n), C.byref(lam)))
    return Y, lam.value


def belos_solve_cheby(rowptr, colind, values, B, cheby, X0=None, block_size=1, num_blocks=300, max_iters=-1,
                      max_restarts=20, tol=1e-8, ortho="DGKS", solver="BlockGmres"):
    """The reference's solver manager with Ifpack_Chebyshev as right preconditioner (Belos::EpetraPrecOp,
    LinearProblem::setRightPrec).  cheby: dict(degree=, lambda_max=, eig_ratio=, min_diag=, eig_iters=, seed=)."""
    rp, ci, va = _csr(rowptr, colind, values)
    B = _F(B)
    n, k = B.shape
    X = np.zeros((n, k), order="F") if X0 is None else _F(X0).copy(order="F")
    o = ortho.encode()
    p = GmresParams(block_size, num_blocks, max_iters, max_restarts, tol, o, 0, -1, 1, 0, 0)
    cp = ChebyParams(cheby.get("degree", 3), cheby.get("lambda_max", -1.0), cheby.get("eig_ratio", 30.0),
                     cheby.get("min_diag", 0.0), cheby.get("eig_iters", 10), cheby.get("seed", 1))
    r = GmresResult()
    _ck(lib().ref_belos_solve_cheby_csr(solver.encode(), C.c_int64(n), _i32(rp), _i32(ci), _d(va), C.byref(cp), k, _d(B),
                                        C.c_int64(n), _d(X), C.c_int64(n), C.byref(p), C.byref(r)))
    return X, dict(converged=bool(r.converged), num_iters=r.num_iters, achieved_tol=r.achieved_tol,
                   loss_of_accuracy=bool(r.loss_of_accuracy), solve_seconds=r.solve_seconds,
                   true_rel_resid=[r.true_rel_resid[i] for i in range(min(k, 64))])


def read_matrix(path, kind="hb"):
    """The reference's own reader: kind "hb" = Trilinos_Util_read_hb + the row insertion of
    Belos::Util::createEpetraProblem; kind "mm" = Trilinos_Util_ReadMatrixMarket2Epetra.
    Returns (n, rowptr, colind, values, b, xexact) (b, xexact only for "hb")."""
    k = 0 if kind == "hb" else 1
    n, nnz = C.c_int64(), C.c_int64()
    pth = os.fsencode(path)
    _ck(lib().ref_read_matrix(k, pth, C.byref(n), C.byref(nnz), None, None, None, None, None))
    rp = np.zeros(n.value + 1, dtype=np.int32)
    ci = np.zeros(max(nnz.value, 1), dtype=np.int32)
    va = np.zeros(max(nnz.value, 1))
    b = np.zeros(max(n.value, 1))
    xe = np.zeros(max(n.value, 1))
    _ck(lib().ref_read_matrix(k, pth, C.byref(n), C.byref(nnz), _i32(rp), _i32(ci), _d(va), _d(b), _d(xe)))
    return n.value, rp, ci[:nnz.value], va[:nnz.value], b[:n.value], xe[:n.value]


def timer_summary():
    buf = C.create_string_buffer(1 << 16)
    lib().ref_timer_summary(buf, len(buf))
    return buf.value.decode(errors="replace")
