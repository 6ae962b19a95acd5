"""Python mirror of the Belos plugin surface over the C-ABI (tests / bench harness).

Method names, argument meaning and error behaviour follow Belos::MultiVec<double> /
Belos::Operator<double> (belos/src/BelosMultiVec.hpp:90-294, BelosOperator.hpp:79-145) as
implemented by the reference adapter (belos/epetra/src/BelosEpetraAdapter.cpp:318-640), so the
parity tests read like the reference's own MVOPTester.  numpy is used only to hold host
buffers; all arithmetic on n-sized data happens on the GPU.
"""
import ctypes as C
import os

import numpy as np

from ._lib import B200Error, check, lib, c_i32, c_i64, c_dbl, vp, P

NORM_ONE, NORM_TWO, NORM_INF = 0, 1, 2
KCLASS = ("spmm", "trans_mv", "times_mat_add_mv", "dot_norm", "axpby_scale_copy", "halo",
          "times_mat_add_mv_with_next_trans_mv", "findbasis_column_step")


def _dptr(a):
    return a.ctypes.data_as(P(c_dbl))


def _iptr(a):
    return a.ctypes.data_as(P(C.c_int))


class Context:
    """One per process / GPU (replaces Epetra_Comm)."""

    def __init__(self, device=0):
        h = vp()
        check(lib.b200_ctx_create(device, C.byref(h)))
        self._h = h
        self.device = device

    def close(self):
        if self._h:
            lib.b200_ctx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def sync(self):
        check(lib.b200_ctx_sync(self._h))

    def set_seed(self, seed):
        check(lib.b200_ctx_set_seed(self._h, int(seed)))

    def comm_init(self, unique_id: bytes, nranks: int, rank: int):
        buf = C.create_string_buffer(unique_id, 128)
        check(lib.b200_ctx_comm_init(self._h, buf, nranks, rank))

    @staticmethod
    def unique_id() -> bytes:
        buf = C.create_string_buffer(128)
        check(lib.b200_comm_unique_id(buf))
        return buf.raw

    @property
    def rank(self):
        r, n = C.c_int(), C.c_int()
        check(lib.b200_ctx_rank(self._h, C.byref(r), C.byref(n)))
        return r.value, n.value

    def counters(self):
        a, b = c_i64(), c_i64()
        check(lib.b200_ctx_counters(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def set_fusion(self, on=True):
        check(lib.b200_ctx_set_fusion(self._h, 1 if on else 0))

    def dot_hits(self):
        a = c_i64()
        check(lib.b200_ctx_fusion_stats(self._h, C.byref(a)))
        return a.value

    def spec_gram_stats(self):
        """(updates that ran fused with the next Gram, MvTransMv calls answered from the mailbox)"""
        a, b = c_i64(), c_i64()
        check(lib.b200_ctx_spec_gram_stats(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def answered_bytes(self, reset=True):
        """SURVEY 8(d) bytes of the reduction calls answered from the memos (no kernel of their own)"""
        a = c_dbl()
        check(lib.b200_ctx_answered_bytes(self._h, C.byref(a), 1 if reset else 0))
        return a.value

    def findbasis_stats(self):
        """one-column MvTransMv calls answered from remembered products (no launch)"""
        a = c_i64()
        check(lib.b200_ctx_findbasis_stats(self._h, C.byref(a)))
        return a.value

    def deferral_stats(self):
        """(lazy single-column copies never launched, single-column scales folded into the next skinny Gram)"""
        a, b = c_i64(), c_i64()
        check(lib.b200_ctx_deferral_stats(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def timing_enable(self, on=True):
        check(lib.b200_ctx_timing_enable(self._h, 1 if on else 0))

    def timing_read(self, reset=True):
        ms = (c_dbl * len(KCLASS))()
        calls = (c_i64 * len(KCLASS))()
        nbytes = (c_dbl * len(KCLASS))()
        check(lib.b200_ctx_timing_read(self._h, ms, calls, nbytes, 1 if reset else 0))
        return {k: {"ms": ms[i], "calls": calls[i], "bytes": nbytes[i]} for i, k in enumerate(KCLASS)}

    def flush_l2(self):
        check(lib.b200_ctx_flush_l2(self._h))

    @property
    def stream(self):
        return lib.b200_ctx_stream(self._h)


class MultiVec:
    """Belos::MultiVec<double> over device storage (column-major, FP64)."""

    def __init__(self, ctx, n_local, ncols, n_global=None, row_offset=0, zero=True, _h=None, _keep=None):
        self.ctx = ctx
        self._keep = _keep  # parent kept alive for views (storage is ref-counted in C as well)
        if _h is not None:
            self._h = _h
            return
        h = vp()
        ng = n_local if n_global is None else n_global
        check(lib.b200_mv_create(ctx._h, n_local, ng, row_offset, ncols, 1 if zero else 0, C.byref(h)))
        self._h = h

    def Multiply(self, scalar_ab, A, B, scalar_this):
        """Epetra_MultiVector::Multiply(ScalarAB, A, B, ScalarThis), element-wise, A a single vector."""
        check(lib.b200_mv_multiply_elementwise(self._h, scalar_ab, A._h, B._h, scalar_this))

    @classmethod
    def from_numpy(cls, ctx, a, n_global=None, row_offset=0):
        a = np.asfortranarray(np.atleast_2d(a.T).T if a.ndim == 1 else a, dtype=np.float64)
        mv = cls(ctx, a.shape[0], a.shape[1], n_global, row_offset, zero=False)
        mv.set_numpy(a)
        return mv

    def __del__(self):
        try:
            # (after the context was closed its objects are gone with it: nothing to release, and nothing safe to touch)
            if self._h and getattr(self.ctx, "_h", None):
                lib.b200_mv_destroy(self._h)
                self._h = None
        except Exception:
            pass

    # --- Belos::MultiVec interface -------------------------------------------------
    def Clone(self, numvecs):
        h = vp()
        check(lib.b200_mv_clone(self._h, numvecs, C.byref(h)))
        return MultiVec(self.ctx, 0, 0, _h=h)

    def CloneCopy(self, index=None):
        h = vp()
        if index is None:
            check(lib.b200_mv_clone_copy(self._h, None, 0, C.byref(h)))
        else:
            idx = np.ascontiguousarray(index, dtype=np.intc)
            check(lib.b200_mv_clone_copy(self._h, _iptr(idx), len(idx), C.byref(h)))
        return MultiVec(self.ctx, 0, 0, _h=h)

    def CloneViewNonConst(self, index):
        idx = np.ascontiguousarray(index, dtype=np.intc)
        h = vp()
        check(lib.b200_mv_clone_view(self._h, _iptr(idx), len(idx), C.byref(h)))
        return MultiVec(self.ctx, 0, 0, _h=h, _keep=self)

    CloneView = CloneViewNonConst

    def GetGlobalLength(self):
        return lib.b200_mv_global_length(self._h)

    def GetLocalLength(self):
        return lib.b200_mv_local_length(self._h)

    def GetNumberVecs(self):
        return lib.b200_mv_num_vecs(self._h)

    def MvTimesMatAddMv(self, alpha, A, B, beta):
        B = np.asarray(B, dtype=np.float64)
        if B.ndim == 1:
            B = B.reshape(-1, 1)
        Bf = np.asfortranarray(B)
        ldb = max(1, Bf.shape[0])
        check(lib.b200_mv_times_mat_add_mv(self._h, alpha, A._h, _dptr(Bf), ldb, Bf.shape[0], Bf.shape[1], beta))

    def MvAddMv(self, alpha, A, beta, B):
        check(lib.b200_mv_add_mv(self._h, alpha, A._h, beta, B._h))

    def MvScale(self, alpha):
        if np.isscalar(alpha):
            check(lib.b200_mv_scale(self._h, float(alpha)))
        else:
            a = np.ascontiguousarray(alpha, dtype=np.float64)
            check(lib.b200_mv_scale_cols(self._h, _dptr(a), len(a)))

    def MvTransMv(self, alpha, A, B=None, ldb=None):
        """B <- alpha * A^T * self.  Returns the p x q block (optionally written into a
        caller-provided Fortran-ordered array with leading dimension ldb)."""
        p, q = A.GetNumberVecs(), self.GetNumberVecs()
        if B is None:
            B = np.zeros((p, q), order="F")
        ld = max(1, B.shape[0]) if ldb is None else ldb
        check(lib.b200_mv_trans_mv(self._h, alpha, A._h, _dptr(B), ld, p, q))
        return B

    def MvDot(self, A):
        k = self.GetNumberVecs()
        out = np.zeros(max(k, 1))
        check(lib.b200_mv_dot(self._h, A._h, _dptr(out), k))
        return out[:k]

    def MvNorm(self, norm_type=NORM_TWO):
        k = self.GetNumberVecs()
        out = np.zeros(max(k, 1))
        check(lib.b200_mv_norm(self._h, norm_type, _dptr(out), k))
        return out[:k]

    def SetBlock(self, A, index):
        idx = np.ascontiguousarray(index, dtype=np.intc)
        check(lib.b200_mv_set_block(self._h, A._h, _iptr(idx), len(idx)))

    def MvRandom(self):
        check(lib.b200_mv_random(self._h))

    def MvInit(self, alpha=0.0):
        check(lib.b200_mv_init(self._h, float(alpha)))

    # --- TSQR hooks (Belos::MultiVec::factorExplicit / revealRank) --------------------
    def factorExplicit(self, Q, force_nonnegative_diagonal=False):
        """self = Q R; Q: multivector with as many columns.  Returns R (b x b)."""
        b = self.GetNumberVecs()
        R = np.zeros((b, b), order="F")
        check(lib.b200_mv_factor_explicit(self._h, Q._h, _dptr(R), max(1, b), 1 if force_nonnegative_diagonal else 0))
        return R

    def revealRank(self, R, tol):
        """self = the Q of factorExplicit, R its factor (modified in place when rank deficient).  Returns the rank."""
        assert R.flags.f_contiguous
        rank = C.c_int(0)
        check(lib.b200_mv_reveal_rank(self._h, _dptr(R), max(1, R.shape[0]), float(tol), C.byref(rank)))
        return rank.value

    # --- host transfer ---------------------------------------------------------------
    def to_numpy(self):
        n, k = self.GetLocalLength(), self.GetNumberVecs()
        out = np.zeros((n, k), order="F")
        if k > 0:
            check(lib.b200_mv_copy_to_host(self._h, _dptr(out), max(n, 1)))
        return out

    def set_numpy(self, a):
        a = np.asfortranarray(a, dtype=np.float64)
        if a.ndim == 1:
            a = a.reshape(-1, 1, order="F")
        n, k = self.GetLocalLength(), self.GetNumberVecs()
        if a.shape != (n, k):
            raise ValueError(f"shape {a.shape} != multivector shape {(n, k)}")
        if k > 0:
            check(lib.b200_mv_copy_from_host(self._h, _dptr(a), max(n, 1)))


class ChebyParams(C.Structure):
    """include/b200_belos.h b200_cheby_params (Ifpack_Chebyshev ParameterList keys)."""
    _fields_ = [("degree", C.c_int), ("lambda_max", c_dbl), ("eig_ratio", c_dbl), ("min_diag", c_dbl),
                ("eig_iters", C.c_int), ("seed", C.c_uint)]


class Chebyshev:
    """Chebyshev smoother in D^-1 A: the reference's Ifpack_Chebyshev (point mode, zero starting solution;
    ifpack/src/Ifpack_Chebyshev.cpp) on the B200 kernels.  Apply == ApplyInverse, so the object is used as a
    Belos::Operator preconditioner (Belos::EpetraPrecOp's role)."""

    def __init__(self, op, degree=1, lambda_max=-1.0, eig_ratio=30.0, min_diag=0.0, eig_iters=10, seed=1):
        self.op = op  # must outlive the smoother
        self.ctx = op.ctx
        self.params = ChebyParams(degree, lambda_max, eig_ratio, min_diag, eig_iters, seed)
        h = vp()
        check(lib.b200_cheby_create(op._h, C.byref(self.params), C.byref(h)))
        self._h = h

    def __del__(self):
        try:
            # (after the context was closed its objects are gone with it: nothing to release, and nothing safe to touch)
            if self._h and getattr(self.ctx, "_h", None):
                lib.b200_cheby_destroy(self._h)
                self._h = None
        except Exception:
            pass

    def ApplyInverse(self, x, y):
        check(lib.b200_cheby_apply(self._h, x._h, y._h))

    Apply = ApplyInverse

    def lambdas(self):
        a, b = c_dbl(), c_dbl()
        check(lib.b200_cheby_lambda(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value


class CrsOperator:
    """Belos::Operator<double> over a row-block CSR matrix (replaces Belos::EpetraOp over
    Epetra_CrsMatrix)."""

    def __init__(self, ctx, n_global, row_begin, row_end, rowptr, colind_global, values):
        self.ctx = ctx
        rp = np.ascontiguousarray(rowptr, dtype=np.int32)
        ci = np.ascontiguousarray(colind_global, dtype=np.int64)
        va = np.ascontiguousarray(values, dtype=np.float64)
        h = vp()
        check(lib.b200_op_create_csr(ctx._h, n_global, row_begin, row_end, rp.ctypes.data_as(P(c_i32)),
                                     ci.ctypes.data_as(P(c_i64)), _dptr(va), C.byref(h)))
        self._h = h
        self.n_global, self.row_begin, self.row_end = n_global, row_begin, row_end

    def __del__(self):
        try:
            # (after the context was closed its objects are gone with it: nothing to release, and nothing safe to touch)
            if self._h and getattr(self.ctx, "_h", None):
                lib.b200_op_destroy(self._h)
                self._h = None
        except Exception:
            pass

    def Apply(self, x, y, trans=0):
        check(lib.b200_op_apply(self._h, x._h, y._h, trans))

    def HasApplyTranspose(self):
        return False

    def ExtractDiagonalCopy(self):
        """Epetra_CrsMatrix::ExtractDiagonalCopy -> one-column MultiVec on the row map."""
        h = vp()
        check(lib.b200_op_extract_diagonal(self._h, C.byref(h)))
        return MultiVec(self.ctx, 0, 1, _h=h)

    def info(self):
        v = [c_i64() for _ in range(5)]
        check(lib.b200_op_info(self._h, *[C.byref(x) for x in v]))
        return dict(zip(("n_local", "n_global", "nnz_local", "n_ghost", "n_boundary_rows"), [x.value for x in v]))


# ---------------------------------------------------------------- host planning helpers
def linear_partition(n_global, nranks, rank):
    a, b = c_i64(), c_i64()
    check(lib.b200_linear_partition(n_global, nranks, rank, C.byref(a), C.byref(b)))
    return a.value, b.value


def build_colmap(n_global, nranks, rank, colind_global):
    ci = np.ascontiguousarray(colind_global, dtype=np.int64)
    nnz = len(ci)
    loc = np.zeros(max(nnz, 1), dtype=np.int32)
    gids = np.zeros(max(nnz, 1), dtype=np.int64)
    own = np.zeros(max(nnz, 1), dtype=np.int32)
    ng = lib.b200_build_colmap(n_global, nranks, rank, nnz, ci.ctypes.data_as(P(c_i64)),
                               loc.ctypes.data_as(P(c_i32)), gids.ctypes.data_as(P(c_i64)),
                               own.ctypes.data_as(P(c_i32)))
    if ng < 0:
        raise B200Error(-1, lib.b200_last_error().decode())
    return loc[:nnz], gids[:ng], own[:ng]


def read_hb(path, row_begin=0, row_end=-1):
    """Harwell-Boeing file -> (n, rowptr, colind_global, values, rhs or None, xexact or None) for rows
    [row_begin, row_end): the matrix Belos::Util::createEpetraProblem builds (b200_read_hb)."""
    n, nnz, has = c_i64(), c_i64(), C.c_int()
    pth = os.fsencode(path)
    check(lib.b200_read_hb(pth, row_begin, row_end, C.byref(n), C.byref(nnz), None, None, None, None, None, C.byref(has)))
    re_ = n.value if row_end < 0 else row_end
    nl = re_ - row_begin
    rp = np.zeros(nl + 1, dtype=np.int32)
    ci = np.zeros(max(nnz.value, 1), dtype=np.int64)
    va = np.zeros(max(nnz.value, 1), dtype=np.float64)
    rhs = np.zeros(max(nl, 1)) if has.value else None
    xe = np.zeros(max(nl, 1)) if has.value else None
    check(lib.b200_read_hb(pth, row_begin, row_end, C.byref(n), C.byref(nnz), rp.ctypes.data_as(P(c_i32)),
                           ci.ctypes.data_as(P(c_i64)), _dptr(va), _dptr(rhs) if has.value else None,
                           _dptr(xe) if has.value else None, C.byref(has)))
    return (n.value, rp, ci[:nnz.value], va[:nnz.value], rhs[:nl] if has.value else None, xe[:nl] if has.value else None)


def read_matrix_market(path, row_begin=0, row_end=-1):
    """MatrixMarket coordinate real general|symmetric -> (n, rowptr, colind_global, values) (b200_read_matrix_market)."""
    n, nnz = c_i64(), c_i64()
    pth = os.fsencode(path)
    check(lib.b200_read_matrix_market(pth, row_begin, row_end, C.byref(n), C.byref(nnz), None, None, None))
    re_ = n.value if row_end < 0 else row_end
    rp = np.zeros(re_ - row_begin + 1, dtype=np.int32)
    ci = np.zeros(max(nnz.value, 1), dtype=np.int64)
    va = np.zeros(max(nnz.value, 1), dtype=np.float64)
    check(lib.b200_read_matrix_market(pth, row_begin, row_end, C.byref(n), C.byref(nnz), rp.ctypes.data_as(P(c_i32)),
                                      ci.ctypes.data_as(P(c_i64)), _dptr(va)))
    return n.value, rp, ci[:nnz.value], va[:nnz.value]


def galeri(name, nx, ny=1, nz=1, conv=1.0, diff=1.0e-5, row_begin=0, row_end=None):
    """Galeri-style stencil matrix rows [row_begin,row_end) in CSR with global column ids."""
    dims = {"Laplace2D": nx * ny, "Recirc2D": nx * ny, "Laplace3D": nx * ny * nz, "Brick3D": nx * ny * nz,
            "Tridiag": nx, "Diag": nx}
    n = dims[name]
    if row_end is None:
        row_end = n
    nnz = c_i64()
    nm = name.encode()
    check(lib.b200_galeri_generate(nm, nx, ny, nz, conv, diff, row_begin, row_end, C.byref(nnz), None, None, None))
    rowptr = np.zeros(row_end - row_begin + 1, dtype=np.int32)
    colind = np.zeros(max(nnz.value, 1), dtype=np.int64)
    values = np.zeros(max(nnz.value, 1), dtype=np.float64)
    check(lib.b200_galeri_generate(nm, nx, ny, nz, conv, diff, row_begin, row_end, C.byref(nnz),
                                   rowptr.ctypes.data_as(P(c_i32)), colind.ctypes.data_as(P(c_i64)),
                                   _dptr(values)))
    return n, rowptr, colind[:nnz.value], values[:nnz.value]
