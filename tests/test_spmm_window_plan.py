"""CPU test (no GPU) of the host side of the window-staged SpMM (trilinos_b200/csrc/spmm_win.cu): the tile plan
(sliced-ELL blobs + x-window runs) is replayed in numpy exactly the way the kernel consumes it -- stage the runs,
then per row sum value * window[index] left to right -- and compared BIT FOR BIT with the oracle's CSR row sums
(Epetra_CrsMatrix::GeneralMM order, epetra/src/Epetra_CrsMatrix.cpp:3516-3534), for serial and row-block
(ghost-column) matrices."""
import ctypes as C

import numpy as np
import pytest

import oracle
import trilinos_b200 as tb
from trilinos_b200._lib import lib

TILE = np.dtype([("blob_off", "<i8"), ("row0", "<i4"), ("nrows", "<i4"), ("width", "<i4"), ("nruns", "<i4"),
                 ("wlen", "<i4"), ("ghost_mask", "<i4"), ("run_start", "<i4", 8), ("run_len", "<i4", 8),
                 ("run_off", "<i4", 8)])
assert TILE.itemsize == 128


def plan(n_local, n_ghost, rp, ci, va, R):
    rp = np.ascontiguousarray(rp, dtype=np.int32)
    ci = np.ascontiguousarray(ci, dtype=np.int32)
    va = np.ascontiguousarray(va, dtype=np.float64)
    nt, bb, bc, wc = C.c_int64(), C.c_int64(), C.c_int32(), C.c_int32()
    args = (n_local, n_ghost, rp.ctypes.data_as(C.POINTER(C.c_int32)), ci.ctypes.data_as(C.POINTER(C.c_int32)),
            va.ctypes.data_as(C.POINTER(C.c_double)), R, C.byref(nt), C.byref(bb), C.byref(bc), C.byref(wc))
    rc = lib.b200_spmm_window_plan(*args, None, None)
    if rc == 1:
        return None
    assert rc == 0
    tiles = np.zeros(nt.value, dtype=TILE)
    blob = np.zeros(bb.value, dtype=np.uint8)
    assert lib.b200_spmm_window_plan(*args, tiles.ctypes.data_as(C.c_void_p), blob.ctypes.data_as(C.c_void_p)) == 0
    return tiles, blob, bc.value, wc.value


def replay(tiles, blob, wcap, R, x_own, x_ghost, n_local):
    """What spmm_win_kernel computes: x_own has ld >= round_up(n_local, 16) rows (padding rows NaN), x_ghost likewise."""
    k = x_own.shape[1]
    y = np.full((n_local, k), np.nan)
    for t in tiles:
        assert t["blob_off"] % 128 == 0 and t["wlen"] <= wcap
        off = int(t["blob_off"])
        row0, nrows, width = np.frombuffer(blob, dtype="<i4", count=3, offset=off)
        assert (row0, nrows, width) == (t["row0"], t["nrows"], t["width"])
        vals = np.frombuffer(blob, dtype="<f8", count=width * R, offset=off + 128).reshape(width, R)
        lcol = np.frombuffer(blob, dtype="<u2", count=width * R, offset=off + 128 + width * R * 8).reshape(width, R)
        cnt = np.frombuffer(blob, dtype="u1", count=R, offset=off + 128 + width * R * 10)
        win = np.full((max(int(t["wlen"]), 1), k), np.nan)
        for j in range(t["nruns"]):
            s, ln, o = int(t["run_start"][j]), int(t["run_len"][j]), int(t["run_off"][j])
            assert s % 16 == 0 and ln % 16 == 0 and ln > 0
            src = x_ghost if (t["ghost_mask"] >> j) & 1 else x_own
            assert s + ln <= src.shape[0], "run reads past the padded column"
            win[o:o + ln] = src[s:s + ln]
        acc = np.zeros((R, k))
        for j in range(width):
            on = j < cnt
            xv = win[lcol[j]]
            acc[on] = acc[on] + vals[j][on, None] * xv[on]     # (numpy: separate multiply and add, like the C++ loop)
        y[row0:row0 + nrows] = acc[:nrows]
        assert np.all(cnt[nrows:] == 0)
    return y


def padded(x, ld):
    out = np.full((ld, x.shape[1]), np.nan)
    out[:x.shape[0]] = x
    return out


@pytest.mark.parametrize("name,dims", [("Laplace3D", (13, 11, 9)), ("Laplace2D", (40, 37, 1)), ("Recirc2D", (300, 7, 1)),
                                       ("Brick3D", (20, 16, 12)), ("Tridiag", (1000, 1, 1))])
@pytest.mark.parametrize("R", [128, 256])
def test_plan_replay_serial(name, dims, R):
    n, rp, ci, va = oracle.galeri_csr(name, *dims)
    p = plan(n, 0, rp, ci, va, R)
    assert p is not None, "stencil matrices must qualify for the window format"
    tiles, blob, bcap, wcap = p
    assert len(tiles) == (n + R - 1) // R
    rng = np.random.default_rng(3)
    x = rng.uniform(-1, 1, (n, 3))
    ld = (n + 31) // 32 * 32
    y = replay(tiles, blob, wcap, R, padded(x, ld), np.zeros((0, 3)), n)
    yref = oracle.spmm(rp, ci.astype(np.int32), va, x)
    assert np.array_equal(y, yref)


@pytest.mark.parametrize("name,dims,P", [("Laplace3D", (12, 10, 16), 3), ("Recirc2D", (50, 40, 1), 4), ("Brick3D", (16, 12, 24), 2)])
def test_plan_replay_row_blocks_with_ghosts(name, dims, P):
    """Row-block partition: ghost columns (local id >= n_local, MakeColMap order) form runs of their own in the halo
    block; tiles split into interior (no ghost run) and boundary; together they reproduce the serial product."""
    n, rp, ci, va = oracle.galeri_csr(name, *dims)
    rng = np.random.default_rng(11)
    x = rng.uniform(-1, 1, (n, 2))
    yref = oracle.spmm(rp, ci.astype(np.int32), va, x)
    seen_bnd = 0
    for r in range(P):
        lo, hi = tb.linear_partition(n, P, r)
        _, rpl, cil, val = tb.galeri(name, *dims, row_begin=lo, row_end=hi)
        loc, gids, own = tb.build_colmap(n, P, r, cil)
        nl, ng = hi - lo, len(gids)
        # rows sorted by local column id, as b200_op_create_csr does
        loc2, val2 = loc.copy(), val.copy()
        for i in range(nl):
            s = slice(rpl[i], rpl[i + 1])
            o = np.argsort(loc[s], kind="stable")
            loc2[s], val2[s] = loc[s][o], val[s][o]
        p = plan(nl, ng, rpl, loc2, val2, 128)
        assert p is not None
        tiles, blob, bcap, wcap = p
        xg = padded(x[gids], (ng + 31) // 32 * 32) if ng else np.zeros((0, 2))
        y = replay(tiles, blob, wcap, 128, padded(x[lo:hi], (nl + 31) // 32 * 32), xg, nl)
        # same entries, sorted by LOCAL id instead of global id: the row sums see them in another order
        yloc = oracle.spmm(rpl, loc2, val2, np.vstack([x[lo:hi], x[gids]]))
        assert np.array_equal(y, yloc)
        assert np.abs(y - yref[lo:hi]).max() <= 1e-13 * np.abs(yref).max()
        seen_bnd += int(np.count_nonzero(tiles["ghost_mask"]))
    assert seen_bnd > 0


def test_format_declines_irregular_matrices():
    """More than 8 x-runs per tile (scattered columns) or heavy padding -> the gather kernel keeps the matrix."""
    rng = np.random.default_rng(5)
    n = 200000
    rp = np.arange(0, 9 * n + 1, 9, dtype=np.int32)
    ci = np.sort(rng.integers(0, n, (n, 9)), axis=1).astype(np.int32).ravel()
    assert plan(n, 0, rp, ci, np.ones(9 * n), 128) is None
    n = 4096
    # one long row among short ones: > 25 % padding
    cnt = np.ones(n, dtype=np.int32)
    cnt[::128] = 64
    rp = np.concatenate([[0], np.cumsum(cnt)]).astype(np.int32)
    ci = np.concatenate([np.arange(i, i + c) % n for i, c in enumerate(cnt)]).astype(np.int32)
    assert plan(n, 0, rp, ci, np.ones(len(ci)), 128) is None
