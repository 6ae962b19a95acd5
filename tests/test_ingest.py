"""matrix ingest.  b200_read_hb / b200_read_matrix_market (host code of libb200belos.so) against the
reference's own readers (Trilinos_Util_read_hb + Belos::Util::createEpetraProblem's row insertion;
Trilinos_Util_ReadMatrixMarket2Epetra), bit-exact: on the generated fixtures of tests/golden (golden CSR produced by the
reference readers, oracle/gen_golden.py --ingest) and, where the reference tree is present, on the reference's own files
(orsirr1.hb of bl_gmres_hb, ...).  GPU part: the files feed the operator and a bl_gmres_hb-style solve."""
import json
import os

import numpy as np
import pytest

import trilinos_b200 as tb
from oracle import ref
from oracle.gen_golden import INGEST

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
K = np.load(os.path.join(GOLD, "ref_ingest.npz"))
SOLVES = json.load(open(os.path.join(GOLD, "ref_solves_ingest.json")))
REF_ROOT = "/root/reference/packages"
REF_HB = ["belos/epetra/example/BlockGmres/orsirr1.hb", "belos/epetra/example/BlockGmres/orsirr1_scaled.hb",
          "belos/test/BlockCG/A.hb", "belos/epetra/test/RCG/gr_30_30.hb", "belos/epetra/test/BiCGStab/cage4.hb",
          "belos/epetra/example/GCRODR/sherman5.hb"]
REF_MM = ["triutils/test/Read/In_bcsstk01.mtx", "triutils/test/Read/In_fs_183_4.mtx", "triutils/test/Read/In_Diagonal.mtx",
          "triutils/test/Read/In_impcol_a.mtx", "belos/epetra/example/BlockGmres/identity.mtx"]


def read(fn, kind, lo=0, hi=-1):
    path = os.path.join(GOLD, fn)
    if kind == "hb":
        return tb.read_hb(path, lo, hi)
    return tb.read_matrix_market(path, lo, hi) + (None, None)


@pytest.mark.parametrize("case", INGEST, ids=[c[0] for c in INGEST])
def test_reader_matches_reference_reader_on_fixtures(case):
    fn, kind, name, dims, sym, with_rhs = case
    n, rp, ci, va, b, xe = read(fn, kind)
    assert np.array_equal(rp, K[fn + "_rowptr"]) and np.array_equal(ci, K[fn + "_colind"])
    assert np.array_equal(va, K[fn + "_values"])            # bit-exact (17 significant digits in the file)
    if with_rhs:
        assert np.array_equal(b, K[fn + "_b"]) and np.array_equal(xe, K[fn + "_xexact"])
    else:
        assert b is None and xe is None
    # row-block read (what a rank of the linear map asks for) == slice of the full read
    lo, hi = n // 3, (2 * n) // 3 + 1
    n2, rp2, ci2, va2, b2, _ = read(fn, kind, lo, hi)
    assert n2 == n and np.array_equal(rp2, rp[lo:hi + 1] - rp[lo])
    assert np.array_equal(ci2, ci[rp[lo]:rp[hi]]) and np.array_equal(va2, va[rp[lo]:rp[hi]])
    if with_rhs:
        assert np.array_equal(b2, b[lo:hi])


def test_reader_errors(tmp_path):
    with pytest.raises(tb.B200Error):
        tb.read_hb(str(tmp_path / "missing.hb"))
    p = tmp_path / "complex.hb"
    p.write_text("t\n 1 1 1 1 0\nCUA 2 2 2 0\n(10I8) (10I8) (3E25.16)\n")
    with pytest.raises(tb.B200Error):
        tb.read_hb(str(p))                                   # the reference: "Can only handle real valued matrices"
    q = tmp_path / "array.mtx"
    q.write_text("%%MatrixMarket matrix array real general\n2 2\n1\n2\n3\n4\n")
    with pytest.raises(tb.B200Error):
        tb.read_matrix_market(str(q))


@pytest.mark.skipif(not (ref.available() and os.path.isdir(REF_ROOT)), reason="needs the reference tree and oracle/_ref")
@pytest.mark.parametrize("rel", REF_HB + REF_MM)
def test_reader_matches_reference_reader_on_reference_files(rel, capfd):
    path = os.path.join(REF_ROOT, rel)
    kind = "hb" if rel.endswith(".hb") else "mm"
    got = tb.read_hb(path)[:4] if kind == "hb" else tb.read_matrix_market(path)
    n2, rp2, ci2, va2, _, _ = ref.read_matrix(path, kind)
    capfd.readouterr()                                       # the reference reader prints a banner
    assert got[0] == n2 and np.array_equal(got[1], rp2) and np.array_equal(got[2], ci2) and np.array_equal(got[3], va2)


@pytest.mark.gpu
@pytest.mark.parametrize("fn", sorted(SOLVES))
def test_solve_from_file(ctx, fn):
    """belos/epetra/test/BlockGmres/test_bl_gmres_hb.cpp in spirit: file -> operator -> Block GMRES, iteration count
    against the reference's run on the matrix its reader produced."""
    import oracle
    g = SOLVES[fn]
    n, rp, ci, va, _, _ = read(fn, g["kind"])
    op = tb.CrsOperator(ctx, n, 0, n, rp, ci, va)
    xs, _ = oracle.random(1, n, 2)
    Xs = tb.MultiVec.from_numpy(ctx, xs)
    B = tb.MultiVec(ctx, n, 2)
    op.Apply(Xs, B)
    X = tb.MultiVec(ctx, n, 2)
    r = tb.BlockGmresSolMgr(op, block_size=2, num_blocks=g["num_blocks"], max_iters=g["max_iters"],
                            max_restarts=g["max_restarts"], tol=g["tol"], ortho="DGKS").solve(X, B)
    assert r["converged"] == g["converged"]
    assert max(r["true_rel_resid"]) <= g["tol"] * 1.001
    assert abs(r["num_iters"] - g["num_iters"]) <= max(1, round(0.02 * g["num_iters"]))
