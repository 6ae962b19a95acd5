"""Generates tests/golden/*.npz from the REFERENCE's own compiled code (oracle/_ref, see
oracle/Makefile.ref).  Run in the build container (needs /root/reference to have built
oracle/_ref); the fixtures are committed and travel to the GPU box.

    python oracle/gen_golden.py            # kernel + small-solve fixtures (seconds)
    python oracle/gen_golden.py --full     # additionally BASELINE config 1/2-sized solves (minutes)
    python oracle/gen_golden.py --configs345   # BASELINE configs 3, 4 (full solves) and 5 (first restart cycle)

Every array below is an output of reference code on explicit, seeded inputs:
  ref_galeri_*.npz   Galeri::CreateCrsMatrix on the "Linear" map                 (CSR, global ids)
  ref_kernels.npz    Belos::MultiVecTraits<double,Epetra_MultiVector> ops, Epetra_CrsMatrix::Multiply,
                     Epetra_MultiVector::Random with explicit seeds
  ref_solves.json    Belos::BlockGmresSolMgr<double,Epetra_MultiVector,Epetra_Operator>: iteration counts,
                     achieved tolerance and true residuals for X_exact = LCG(seed 1), B = A X_exact, X0 = 0
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
from oracle import ref  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")

GALERI_CASES = [("Laplace2D", (10, 8, 1), {}), ("Laplace2D", (100, 100, 1), {}), ("Laplace3D", (7, 6, 5), {}),
                ("Laplace3D", (20, 20, 20), {}), ("Recirc2D", (12, 9, 1), dict(conv=1.0, diff=1e-5)),
                ("Recirc2D", (40, 40, 1), dict(conv=1.0, diff=1e-5))]

# (label, matrix, dims, block size, num blocks, ortho, tol, max restarts[, max iters; default -1 = n/b - 1])
SOLVES_SMALL = [
    ("laplace3d_10_b1_dgks", "Laplace3D", (10, 10, 10), 1, 50, "DGKS", 1e-5, 15),   # the reference's registered example
    ("laplace3d_10_b1_icgs", "Laplace3D", (10, 10, 10), 1, 50, "ICGS", 1e-5, 15),
    ("laplace3d_10_b1_imgs", "Laplace3D", (10, 10, 10), 1, 50, "IMGS", 1e-5, 15),
    ("laplace3d_20_b4_dgks", "Laplace3D", (20, 20, 20), 4, 50, "DGKS", 1e-8, 200),
    ("laplace3d_20_b4_icgs", "Laplace3D", (20, 20, 20), 4, 25, "ICGS", 1e-8, 200),
    ("laplace2d_40_b2_dgks", "Laplace2D", (40, 40, 1), 2, 30, "DGKS", 1e-8, 200),
    ("recirc2d_40_b16_dgks", "Recirc2D", (40, 40, 1), 16, 40, "DGKS", 1e-5, 200, 1000),
    ("brick3d_12_b8_icgs", "Brick3D", (12, 12, 12), 8, 30, "ICGS", 1e-8, 200),
]
# other solver managers on the same kernels: (..., max iters, solver)
SOLVES_MANAGERS = [
    ("pseudo_gmres_laplace3d_12_b4_dgks", "Laplace3D", (12, 12, 12), 4, 40, "DGKS", 1e-8, 50, -1, "PseudoBlockGmres"),
    ("pseudo_gmres_recirc2d_30_b3_icgs", "Recirc2D", (30, 30, 1), 3, 60, "ICGS", 1e-4, 100, -1, "PseudoBlockGmres"),
    ("block_cg_laplace3d_12_b4_dgks", "Laplace3D", (12, 12, 12), 4, 1, "DGKS", 1e-8, 1, -1, "BlockCG"),
    ("block_cg_laplace2d_40_b1_dgks", "Laplace2D", (40, 40, 1), 1, 1, "DGKS", 1e-8, 1, -1, "BlockCG"),
    ("block_cg_brick3d_10_b2_icgs", "Brick3D", (10, 10, 10), 2, 1, "ICGS", 1e-8, 1, -1, "BlockCG"),
    ("pseudo_cg_laplace3d_12_b4", "Laplace3D", (12, 12, 12), 4, 1, "DGKS", 1e-8, 1, -1, "PseudoBlockCG"),
]
SOLVES_FULL = [
    ("config1_laplace2d_100_b1_dgks", "Laplace2D", (100, 100, 1), 1, 50, "DGKS", 1e-8, 1000),
    ("config2_laplace3d_100_b4_dgks", "Laplace3D", (100, 100, 100), 4, 50, "DGKS", 1e-8, 1000),
]


def matrix(name, dims, **kw):
    if name == "Brick3D":  # no Brick3D in Galeri-Epetra: the restated stencil is fed to Epetra as CSR
        return oracle.galeri_csr(name, *dims)
    return ref.galeri(name, *dims, **kw)


def gen_galeri():
    for name, dims, kw in GALERI_CASES:
        n, rp, ci, va = ref.galeri(name, *dims, **kw)
        fn = os.path.join(OUT, f"ref_galeri_{name}_{dims[0]}x{dims[1]}x{dims[2]}.npz")
        np.savez_compressed(fn, n=n, rowptr=rp, colind=ci, values=va, dims=np.array(dims),
                            conv=kw.get("conv", 0.0), diff=kw.get("diff", 0.0))
        print("wrote", fn, n, len(va))


def gen_kernels():
    rng = np.random.default_rng(20261104)
    d = {}
    n, p, q = 1237, 9, 4
    A = np.asfortranarray(rng.uniform(-1, 1, (n, p)))
    V = np.asfortranarray(rng.uniform(-1, 1, (n, q)))
    W = np.asfortranarray(rng.uniform(-1, 1, (n, q)))
    Bm = np.asfortranarray(rng.uniform(-1, 1, (p, q)))
    d.update(A=A, V=V, W=W, Bm=Bm)
    d["trans_mv"] = ref.trans_mv(1.5, A, V)
    d["times_mat_add_mv_m1_1"] = ref.times_mat_add_mv(-1.0, A, Bm, 1.0, V)
    d["times_mat_add_mv_1_0"] = ref.times_mat_add_mv(1.0, A, Bm, 0.0, V)
    d["times_mat_add_mv_h_2"] = ref.times_mat_add_mv(0.5, A, Bm, 2.0, V)
    d["dot"] = ref.dot(V, W)
    d["dot_self"] = ref.dot(V, V)
    for t, nm in ((0, "one"), (1, "two"), (2, "inf")):
        d["norm_" + nm] = ref.norm(A, t)
    d["add_mv_gen"] = ref.add_mv(2.5, V, -0.75, W)
    d["add_mv_m1_1"] = ref.add_mv(-1.0, V, 1.0, W)
    d["add_mv_1_0"] = ref.add_mv(1.0, V, 0.0, W)
    d["add_mv_0_1"] = ref.add_mv(0.0, V, 1.0, W)
    sc = np.array([1.0, 0.0, -2.0, 0.5])
    d["scale_cols_alpha"] = sc
    d["scale_cols"] = ref.scale_cols(sc, V)
    for seed in (1, 12345, 2147483646):
        y, s = ref.random(seed, 501, 3)
        d[f"random_{seed}"] = y
        d[f"random_{seed}_next"] = np.int64(s)
    for name, dims, kw in (("Laplace3D", (7, 6, 5), {}), ("Recirc2D", (12, 9, 1), dict(conv=1.0, diff=1e-5)),
                           ("Laplace2D", (10, 8, 1), {})):
        nn, rp, ci, va = ref.galeri(name, *dims, **kw)
        X = np.asfortranarray(rng.uniform(-1, 1, (nn, 5)))
        d[f"apply_{name}_X"] = X
        d[f"apply_{name}_Y"] = ref.apply(rp, ci, va, X)
    fn = os.path.join(OUT, "ref_kernels.npz")
    np.savez_compressed(fn, **d)
    print("wrote", fn)


# Chebyshev smoother / right preconditioner: the reference's Ifpack_Chebyshev
CHEBY_APPLY = [  # (label, matrix, dims, k, degree, lambda_max (-1: power method, seed 1), ratio)
    ("laplace3d_7x6x5_k1_d3", "Laplace3D", (7, 6, 5), 1, 3, -1.0, 30.0),
    ("laplace3d_7x6x5_k3_d5", "Laplace3D", (7, 6, 5), 3, 5, -1.0, 30.0),
    ("laplace2d_10x8_k4_d1", "Laplace2D", (10, 8, 1), 4, 1, 1.9, 20.0),
    ("recirc2d_12x9_k2_d4", "Recirc2D", (12, 9, 1), 2, 4, 2.5, 30.0),
    ("laplace3d_20_k4_d8", "Laplace3D", (20, 20, 20), 4, 8, -1.0, 30.0),
]
# (label, matrix, dims, block size, num blocks, ortho, tol, max restarts, solver, cheby dict)
SOLVES_PREC = [
    ("gmres_laplace3d_12_b4_cheb4", "Laplace3D", (12, 12, 12), 4, 40, "DGKS", 1e-8, 50, "BlockGmres", dict(degree=4)),
    ("gmres_laplace3d_20_b4_cheb8", "Laplace3D", (20, 20, 20), 4, 50, "DGKS", 1e-8, 50, "BlockGmres", dict(degree=8)),
    ("gmres_laplace2d_40_b1_cheb3", "Laplace2D", (40, 40, 1), 1, 30, "ICGS", 1e-8, 200, "BlockGmres", dict(degree=3, lambda_max=2.0)),
    ("pseudo_gmres_laplace3d_12_b2_cheb2", "Laplace3D", (12, 12, 12), 2, 40, "DGKS", 1e-8, 50, "PseudoBlockGmres", dict(degree=2)),
    ("block_cg_laplace3d_12_b4_cheb3", "Laplace3D", (12, 12, 12), 4, 1, "DGKS", 1e-8, 1, "BlockCG", dict(degree=3)),
]


def gen_cheby():
    out = {}
    for label, name, dims, k, degree, lam, ratio in CHEBY_APPLY:
        kw = dict(conv=1.0, diff=1e-5) if name == "Recirc2D" else {}
        n, rp, ci, va = matrix(name, dims, **kw)
        X, _ = ref.random(7, n, k)
        Y, lam_used = ref.cheby_apply(rp, ci, va, X, degree=degree, lambda_max=lam, eig_ratio=ratio, seed=1)
        out[label + "_X"] = X
        out[label + "_Y"] = Y
        out[label + "_lambda"] = lam_used
        print(label, lam_used, float(np.linalg.norm(Y)))
    np.savez_compressed(os.path.join(OUT, "ref_cheby.npz"), **out)
    path = os.path.join(OUT, "ref_solves_prec.json")
    res = {}
    for label, name, dims, b, m, ortho, tol, maxr, solver, ch in SOLVES_PREC:
        n, rp, ci, va = matrix(name, dims)
        xe, _ = ref.random(1, n, b)
        B = ref.apply(rp, ci, va, xe)
        X, r = ref.belos_solve_cheby(rp, ci, va, B, ch, block_size=b, num_blocks=m, max_restarts=maxr, tol=tol, ortho=ortho,
                                     solver=solver)
        r.update(matrix=name, dims=list(dims), block_size=b, num_blocks=m, ortho=ortho, tol=tol, max_restarts=maxr, max_iters=-1,
                 solver=solver, cheby=ch, max_abs_err=float(np.abs(X - xe).max()),
                 rhs="X_exact = Epetra LCG seed 1, B = A X_exact, X0 = 0; right preconditioner Ifpack_Chebyshev")
        res[label] = r
        print(label, r["num_iters"], r["converged"], max(r["true_rel_resid"]), flush=True)
    json.dump(res, open(path, "w"), indent=1, sort_keys=True)


# matrix ingest: generated files, read by the reference's own readers
INGEST = [  # (file, kind, matrix, dims, symmetric, with rhs)
    ("ingest_recirc2d_12x9.hb", "hb", "Recirc2D", (12, 9, 1), False, True),
    ("ingest_laplace2d_10x8_rsa.hb", "hb", "Laplace2D", (10, 8, 1), True, False),
    ("ingest_laplace3d_7x6x5.mtx", "mm", "Laplace3D", (7, 6, 5), False, False),
    ("ingest_laplace2d_10x8_sym.mtx", "mm", "Laplace2D", (10, 8, 1), True, False),
]


def gen_ingest():
    from oracle import hb_write
    out, solves = {}, {}
    for fn, kind, name, dims, sym, with_rhs in INGEST:
        kw = dict(conv=1.0, diff=1e-5) if name == "Recirc2D" else {}
        n, rp, ci, va = matrix(name, dims, **kw)
        path = os.path.join(OUT, fn)
        if kind == "hb":
            xe = b = None
            if with_rhs:
                xe, _ = ref.random(3, n, 1)
                b = ref.apply(rp, ci, va, xe)[:, 0]
                xe = xe[:, 0]
            hb_write.write_hb(path, n, rp, ci, va, symmetric=sym, rhs=b, xexact=xe)
        else:
            hb_write.write_mm(path, n, rp, ci, va, symmetric=sym)
        n2, rp2, ci2, va2, b2, x2 = ref.read_matrix(path, kind)         # the reference's reader
        assert n2 == n
        out[fn + "_rowptr"], out[fn + "_colind"], out[fn + "_values"] = rp2, ci2, va2
        if with_rhs:
            out[fn + "_b"], out[fn + "_xexact"] = b2, x2
        # bl_gmres_hb-style solve on the matrix as read (belos/epetra/test/BlockGmres/test_bl_gmres_hb.cpp)
        xs, _ = ref.random(1, n, 2)
        B = ref.apply(rp2, ci2, va2, xs)
        X, r = ref.block_gmres(rp2, ci2, va2, B, block_size=2, num_blocks=60, max_iters=200, max_restarts=100, tol=1e-6)
        r.update(file=fn, kind=kind, block_size=2, num_blocks=60, max_iters=200, max_restarts=100, tol=1e-6, ortho="DGKS")
        solves[fn] = r
        print(fn, n, len(va2), r["num_iters"], r["converged"], flush=True)
    np.savez_compressed(os.path.join(OUT, "ref_ingest.npz"), **out)
    json.dump(solves, open(os.path.join(OUT, "ref_solves_ingest.json"), "w"), indent=1, sort_keys=True)


def gen_solves(cases, fn, threads=1):
    path = os.path.join(OUT, fn)
    res = json.load(open(path)) if os.path.exists(path) else {}
    ref.set_threads(threads)
    for case in cases:
        label, name, dims, b, m, ortho, tol, maxr = case[:8]
        max_iters = case[8] if len(case) > 8 else -1
        solver = case[9] if len(case) > 9 else "BlockGmres"
        kw = dict(conv=1.0, diff=1e-5) if name == "Recirc2D" else {}
        n, rp, ci, va = matrix(name, dims, **kw)
        xe, _ = ref.random(1, n, b)
        B = ref.apply(rp, ci, va, xe)
        X, r = ref.block_gmres(rp, ci, va, B, block_size=b, num_blocks=m, max_iters=max_iters, max_restarts=maxr, tol=tol,
                               ortho=ortho, solver=solver)
        r.update(matrix=name, dims=list(dims), block_size=b, num_blocks=m, ortho=ortho, tol=tol, max_restarts=maxr, max_iters=max_iters, solver=solver,
                 rhs="X_exact = Epetra LCG seed 1 (column-major sweep), B = A X_exact, X0 = 0",
                 max_abs_err=float(np.abs(X - xe).max()), threads=threads)
        res[label] = r
        print(label, r["num_iters"], r["converged"], f"{r['solve_seconds']:.2f}s", max(r["true_rel_resid"]), flush=True)
        json.dump(res, open(path, "w"), indent=1, sort_keys=True)
    ref.set_threads(1)


# BASELINE configs 3-5 at full size, all host threads (Epetra OpenMP loops + threaded OpenBLAS), with the native
# residual norms of every block iteration (ref.block_gmres_history).  Config 5 (Recirc2D 2000^2, b=16, m=100) is a
# BOUNDED run: the first 32 block iterations of its first restart cycle.  A single-rank Epetra cannot hold the basis
# of "Num Blocks" = 100 at all: Epetra_MultiVector allocates `new double[Stride_ * NumVectors_]` with int operands
# (epetra/src/Epetra_MultiVector.cpp:257), and 4,000,000 x 1616 overflows it (std::bad_alloc; the reference needs >= 4
# MPI ranks for this config).  The largest basis that fits is 33 blocks (528 columns, 2.11e9 entries), so the run uses
# "Num Blocks" = 32 with "Maximum Iterations" = 32 -- block Arnoldi does not depend on the restart length before the
# first restart, so these 32 iterations are arithmetically identical to the first 32 of the m = 100 cycle (the GPU
# test runs m = 100).
# (label, matrix, dims, block size, num blocks, ortho, tol, max restarts, max iters)
SOLVES_345 = [
    ("config3_laplace3d_200_b4_dgks", "Laplace3D", (200, 200, 200), 4, 50, "DGKS", 1e-8, 1000, -1),
    ("config4_brick3d_150_b8_icgs", "Brick3D", (150, 150, 150), 8, 50, "ICGS", 1e-8, 1000, -1),
    ("config5_recirc2d_2000_b16_dgks_first32", "Recirc2D", (2000, 2000, 1), 16, 32, "DGKS", 1e-8, 0, 32),
]


def gen_345(only=None):
    path = os.path.join(OUT, "ref_solves_full.json")
    res = json.load(open(path)) if os.path.exists(path) else {}
    threads = ref.max_threads()
    ref.set_threads(threads)
    for label, name, dims, b, m, ortho, tol, maxr, max_iters in SOLVES_345:
        if only and only not in label:
            continue
        kw = dict(conv=1.0, diff=1e-5) if name == "Recirc2D" else {}
        n, rp, ci, va = matrix(name, dims, **kw)
        xe, _ = ref.random(1, n, b)
        B = ref.apply(rp, ci, va, xe)
        X, r, hist = ref.block_gmres_history(rp, ci, va, B, block_size=b, num_blocks=m, max_iters=max_iters, max_restarts=maxr,
                                             tol=tol, ortho=ortho)
        if "config5" in label:  # the problem as BASELINE.json states it; what was run is recorded next to it
            r.update(ran_num_blocks=m, ran_max_iters=max_iters)
            m, max_iters, maxr = 100, 100, 0
        r.update(matrix=name, dims=list(dims), block_size=b, num_blocks=m, ortho=ortho, tol=tol, max_restarts=maxr,
                 max_iters=max_iters, solver="BlockGmres",
                 rhs="X_exact = Epetra LCG seed 1 (column-major sweep), B = A X_exact, X0 = 0",
                 max_abs_err=float(np.abs(X - xe).max()), threads=threads,
                 native_resid_history=[[float(v) for v in row] for row in hist],
                 history="row i = Iteration::getNativeResiduals before block iteration i (unscaled 2-norms per right-hand side)")
        res[label] = r
        print(label, r["num_iters"], r["converged"], f"{r['solve_seconds']:.2f}s", max(r["true_rel_resid"]), hist.shape, flush=True)
        json.dump(res, open(path, "w"), indent=1, sort_keys=True)
        del X, B, xe, rp, ci, va
    ref.set_threads(1)


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    if "--ingest" in sys.argv:
        gen_ingest()
    elif "--cheby" in sys.argv:
        gen_cheby()
    elif "--managers" in sys.argv:
        gen_solves(SOLVES_MANAGERS, "ref_solves_managers.json")
    elif "--full" in sys.argv:
        gen_solves(SOLVES_FULL, "ref_solves_full.json", threads=1)
    elif "--configs345" in sys.argv:  # optional: a substring selecting one case
        i = sys.argv.index("--configs345")
        gen_345(sys.argv[i + 1] if len(sys.argv) > i + 1 else None)
    else:
        gen_galeri()
        gen_kernels()
        gen_solves(SOLVES_SMALL, "ref_solves.json")
