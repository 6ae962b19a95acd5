"""CPU tests (no GPU): the C-ABI library loads and exports every symbol include/*.h declares; compute
entry points refuse to run without a device (no CPU fallback); host-side planning (linear map, column
map, halo plan) matches the numpy restatement of Epetra_BlockMap / Epetra_CrsGraph::MakeColMap."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import oracle
import trilinos_b200 as tb
from trilinos_b200._lib import lib
from trilinos_b200.solver import SOLVER_EXPORTED, solver_lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared(header):
    txt = open(os.path.join(ROOT, "include", header)).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", txt)))


def test_every_declared_symbol_is_exported():
    names = declared("b200_belos.h")
    assert len(names) >= 40
    for nm in names:
        assert hasattr(lib, nm), f"{nm} declared in include/b200_belos.h but not exported"
    assert set(names) == set(tb.EXPORTED), set(names) ^ set(tb.EXPORTED)


def test_solver_library_exports():
    sl = solver_lib()
    names = declared("b200_belos_solver.h")
    for nm in names:
        assert hasattr(sl, nm), nm
    assert set(names) == set(SOLVER_EXPORTED)


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(tb.B200Error) as e:
        tb.Context(0)
    assert e.value.code == -4  # B200_ERR_NO_DEVICE
    assert "no CPU fallback" in str(e.value)


def test_product_does_not_import_oracle():
    """The product path must never route through the oracle (test infrastructure)."""
    for dp, _, files in os.walk(os.path.join(ROOT, "trilinos_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp")) or f == "Makefile":
                txt = open(os.path.join(dp, f), errors="replace").read()
                assert not re.search(r"^\s*(import|from)\s+oracle", txt, flags=re.M), os.path.join(dp, f)
                assert "liboracle" not in txt and "libref_epetra" not in txt, os.path.join(dp, f)


@pytest.mark.parametrize("n,P", [(10, 1), (10, 3), (7, 8), (1000003, 8), (0, 2), (5, 5)])
def test_linear_partition(n, P):
    """Epetra_Map(n, 0, comm): first n mod P ranks own ceil(n/P) rows (Epetra_BlockMap.cpp:83-90)."""
    prev = 0
    for r in range(P):
        lo, hi = tb.linear_partition(n, P, r)
        assert (lo, hi) == oracle.linear_partition(n, P, r)
        assert lo == prev and hi - lo in (n // P, n // P + 1)
        prev = hi
    assert prev == n


@pytest.mark.parametrize("name,dims,P", [("Laplace3D", (6, 5, 7), 2), ("Laplace3D", (6, 5, 7), 5), ("Laplace2D", (13, 11, 1), 3),
                                         ("Brick3D", (5, 6, 8), 4), ("Recirc2D", (9, 14, 1), 8), ("Tridiag", (37, 1, 1), 7)])
def test_colmap_and_halo_plan(name, dims, P):
    """MakeColMap ordering (owned first, ghosts grouped by owner, ascending GID) and the Import plan:
    simulating the exchange on the host reproduces the serial SpMM bit for bit."""
    n, rp, ci, va = oracle.galeri_csr(name, *dims)
    rng = np.random.default_rng(5)
    x = rng.uniform(-1, 1, (n, 3))
    y_serial = oracle.spmm(rp, ci.astype(np.int32), va, x)
    parts = []
    for r in range(P):
        lo, hi = tb.linear_partition(n, P, r)
        n2, rpl, cil, val = tb.galeri(name, *dims, row_begin=lo, row_end=hi)
        loc, gids, own = tb.build_colmap(n, P, r, cil)
        loc_o, gids_o, own_o = oracle.make_colmap(n, P, r, cil)
        assert np.array_equal(loc, loc_o) and np.array_equal(gids, gids_o) and np.array_equal(own, own_o)
        need = np.zeros(P, dtype=np.int64)
        off = np.zeros(P, dtype=np.int64)
        own_c = np.ascontiguousarray(own, dtype=np.int32)
        rc = lib.b200_halo_recv_plan(P, len(gids), own_c.ctypes.data_as(C.POINTER(C.c_int32)),
                                     need.ctypes.data_as(C.POINTER(C.c_int64)), off.ctypes.data_as(C.POINTER(C.c_int64)))
        assert rc == 0 and need.sum() == len(gids) and need[r] == 0
        parts.append(dict(lo=lo, hi=hi, rp=rpl, loc=loc, va=val, gids=gids, need=need, off=off))
    for r, pr in enumerate(parts):
        # what the peers ask of me, concatenated by peer rank
        asked = np.concatenate([parts[s]["gids"][parts[s]["off"][r]:parts[s]["off"][r] + parts[s]["need"][r]]
                                for s in range(P)] + [np.zeros(0, dtype=np.int64)])
        sidx = np.zeros(max(len(asked), 1), dtype=np.int32)
        rc = lib.b200_halo_send_index(n, P, r, len(asked), np.ascontiguousarray(asked).ctypes.data_as(C.POINTER(C.c_int64)),
                                      sidx.ctypes.data_as(C.POINTER(C.c_int32)))
        assert rc == 0
        assert np.array_equal(sidx[:len(asked)] + pr["lo"], asked)
        pr["sent"] = x[pr["lo"] + sidx[:len(asked)]]
    for r, pr in enumerate(parts):
        # my ghost tail = what each owner sent me, in owner order
        ghosts = x[pr["gids"]] if len(pr["gids"]) else np.zeros((0, 3))
        xcol = np.vstack([x[pr["lo"]:pr["hi"]], ghosts])
        y = oracle.spmm(pr["rp"], pr["loc"], pr["va"], xcol)
        assert np.array_equal(y, y_serial[pr["lo"]:pr["hi"]])
    # a rank asked for a row it does not own -> error, not silent garbage
    bad = np.array([n + 5], dtype=np.int64)
    out = np.zeros(1, dtype=np.int32)
    assert lib.b200_halo_send_index(n, P, 0, 1, bad.ctypes.data_as(C.POINTER(C.c_int64)),
                                    out.ctypes.data_as(C.POINTER(C.c_int32))) != 0
