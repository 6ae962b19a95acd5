"""numpy restatement of the Galeri generators, the linear map and the column map.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Written independently of the product's
host generator (trilinos_b200/csrc/host_plan.cpp) — vectorised COO assembly instead of a row
loop — so the two can be checked against each other and against the golden fixtures produced
by the reference's compiled Galeri (tests/golden/ref_galeri_*.npz).

Reference: galeri/src-epetra/Galeri_CrsMatrices.cpp:273-288 (Laplace2D), :348-368 (Recirc2D),
:398-416 (Laplace3D); CrsMatrices/Galeri_Cross2D.h:57-117, Galeri_Cross3D.h:53-131,
Galeri_Recirc2D.h:55-124; neighbours Galeri_Utils.cpp:330-398;
Brick3D galeri/src-xpetra/Galeri_XpetraMatrixTypes.hpp:531-633,736-761 with the defaults of
Galeri_StencilProblems.hpp:253-280 and Galeri_Problem.hpp:108-118.
Maps: epetra/src/Epetra_BlockMap.cpp:83-90; column map epetra/src/Epetra_CrsGraph.cpp:1348-1524.
"""
import numpy as np
import scipy.sparse as sp


def linear_partition(n, nranks, rank):
    q, r = divmod(n, nranks)
    lo = rank * q + min(rank, r)
    return lo, lo + q + (1 if rank < r else 0)


def _csr(n, rows, cols, vals, row_begin, row_end):
    A = sp.coo_matrix((vals, (rows, cols)), shape=(n, n)).tocsr()
    A.sum_duplicates()
    A.sort_indices()
    A = A[row_begin:row_end]
    return A.indptr.astype(np.int32), A.indices.astype(np.int64), A.data.astype(np.float64)


def galeri_csr(name, nx, ny=1, nz=1, conv=1.0, diff=1.0e-5, row_begin=0, row_end=None):
    """rows [row_begin,row_end) in CSR, global column ids ascending inside each row."""
    if name in ("Laplace2D", "Recirc2D"):
        n = nx * ny
        g = np.arange(n, dtype=np.int64)
        ix, iy = g % nx, g // nx
        if name == "Laplace2D":
            a = np.full(n, 4.0)
            b = c = d = e = np.full(n, -1.0)
        else:
            hx, hy = 1.0 / (nx + 1), 1.0 / (ny + 1)
            x, y = hx * (ix + 1), hy * (iy + 1)
            cx = conv * 4 * x * (x - 1.0) * (1.0 - 2 * y) / hx
            cy = -conv * 4 * y * (y - 1.0) * (1.0 - 2 * x) / hy
            a = np.zeros(n); b = np.zeros(n); c = np.zeros(n); d = np.zeros(n); e = np.zeros(n)
            neg = cx < 0
            c[neg] += cx[neg]; a[neg] -= cx[neg]
            b[~neg] -= cx[~neg]; a[~neg] += cx[~neg]
            neg = cy < 0
            e[neg] += cy[neg]; a[neg] -= cy[neg]
            d[~neg] -= cy[~neg]; a[~neg] += cy[~neg]
            a = a + (diff * 2.0 / (hx * hx) + diff * 2.0 / (hy * hy))
            b = b - diff / (hx * hx)
            c = c - diff / (hx * hx)
            d = d - diff / (hy * hy)
            e = e - diff / (hy * hy)
        R, Cc, V = [g], [g], [a]
        for mask, off, val in ((ix > 0, -1, b), (ix < nx - 1, 1, c), (iy > 0, -nx, d), (iy < ny - 1, nx, e)):
            R.append(g[mask]); Cc.append(g[mask] + off); V.append(val[mask])
    elif name == "Laplace3D":
        n = nx * ny * nz
        g = np.arange(n, dtype=np.int64)
        ix, iy, iz = g % nx, (g // nx) % ny, g // (nx * ny)
        R, Cc, V = [g], [g], [np.full(n, 6.0)]
        for mask, off in ((ix > 0, -1), (ix < nx - 1, 1), (iy > 0, -nx), (iy < ny - 1, nx),
                          (iz > 0, -nx * ny), (iz < nz - 1, nx * ny)):
            R.append(g[mask]); Cc.append(g[mask] + off); V.append(np.full(mask.sum(), -1.0))
    elif name == "Brick3D":
        n = nx * ny * nz
        g = np.arange(n, dtype=np.int64)
        ix, iy, iz = g % nx, (g // nx) % ny, g // (nx * ny)
        R, Cc, V = [], [], []
        cnt = np.zeros(n)
        for dz in (-1, 0, 1):
            for dy in (-1, 0, 1):
                for dx in (-1, 0, 1):
                    if dx == dy == dz == 0:
                        continue
                    mask = ((ix + dx >= 0) & (ix + dx < nx) & (iy + dy >= 0) & (iy + dy < ny) &
                            (iz + dz >= 0) & (iz + dz < nz))
                    R.append(g[mask]); Cc.append(g[mask] + dx + dy * nx + dz * nx * ny)
                    V.append(np.full(mask.sum(), -1.0))
                    cnt += mask
        xy_bnd = (ix == 0) | (ix == nx - 1) | (iy == 0) | (iy == ny - 1)   # "Dirichlet" test (x/y faces only)
        z_bnd = (iz == 0) | (iz == nz - 1)
        diag = np.where(z_bnd & ~xy_bnd, cnt, 26.0)                         # "Neumann" diagonal on pure z faces
        R.append(g); Cc.append(g); V.append(diag)
    elif name == "Tridiag":
        n = nx
        g = np.arange(n, dtype=np.int64)
        R, Cc, V = [g, g[1:], g[:-1]], [g, g[1:] - 1, g[:-1] + 1], [np.full(n, 2.0), np.full(n - 1, -1.0), np.full(n - 1, -1.0)]
    elif name == "Diag":
        n = nx
        g = np.arange(n, dtype=np.int64)
        R, Cc, V = [g], [g], [(g + 1.0) * (1.0 if conv == 0.0 else conv)]
    else:
        raise ValueError(name)
    if row_end is None:
        row_end = n
    rp, ci, va = _csr(n, np.concatenate(R), np.concatenate(Cc), np.concatenate(V), row_begin, row_end)
    return n, rp, ci, va


def make_colmap(n_global, nranks, rank, colind_global):
    """local column ids: owned GIDs first (domain order), then remote GIDs grouped by owner
    rank, ascending GID inside an owner.  Returns (colind_local, ghost_gids, ghost_owner)."""
    rb, re = linear_partition(n_global, nranks, rank)
    ci = np.asarray(colind_global, dtype=np.int64)
    remote = np.unique(ci[(ci < rb) | (ci >= re)])
    bounds = np.array([linear_partition(n_global, nranks, r)[0] for r in range(nranks)] + [n_global])
    owner = np.searchsorted(bounds, remote, side="right") - 1
    loc = np.where((ci >= rb) & (ci < re), ci - rb, (re - rb) + np.searchsorted(remote, ci))
    return loc.astype(np.int32), remote, owner.astype(np.int32)
