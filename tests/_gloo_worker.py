"""Worker for tests/test_multirank_gloo.py (launched by torch.distributed.run, gloo backend, CPU).

Each rank runs the product's host-side multi-rank logic — linear partition, local rows of the Galeri
matrix, column map, halo plan (C-ABI host functions) — and performs the exchange choreography of
b200_op_apply / the allreduce of the reductions with gloo in place of NCCL; the n-sized arithmetic,
which on the GPU box is done by the CUDA kernels, is done here by the oracle.  Result must equal the
serial oracle: bit for bit for the SpMM, to rounding for the allreduced Gram block.
"""
import ctypes as C
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
import trilinos_b200 as tb  # noqa: E402
from trilinos_b200._lib import lib  # noqa: E402


def main():
    rank, world, _ = tb.dist.env_ranks()
    dist.init_process_group("gloo", rank=rank, world_size=world)
    for name, dims, k in (("Laplace3D", (6, 7, 9), 4), ("Recirc2D", (17, 12, 1), 3), ("Brick3D", (5, 5, 7), 2)):
        n = tb.dist.grid_size(name, dims)
        lo, hi = tb.linear_partition(n, world, rank)
        _, rp, ci, va = tb.galeri(name, *dims, row_begin=lo, row_end=hi)
        loc, gids, own = tb.build_colmap(n, world, rank, ci)
        need = np.zeros(world, dtype=np.int64)
        off = np.zeros(world, dtype=np.int64)
        own32 = np.ascontiguousarray(own, dtype=np.int32)
        assert lib.b200_halo_recv_plan(world, len(gids), own32.ctypes.data_as(C.POINTER(C.c_int32)),
                                       need.ctypes.data_as(C.POINTER(C.c_int64)),
                                       off.ctypes.data_as(C.POINTER(C.c_int64))) == 0
        # "needs" matrix: everyone learns who needs how many rows from whom (allreduce in the product)
        M = torch.zeros((world, world), dtype=torch.int64)
        M[rank] = torch.from_numpy(need)
        dist.all_reduce(M)
        send_cnt = M[:, rank].numpy()
        # exchange the GID lists: send my needs to their owners, receive what is asked of me
        reqs, asked_parts = [], []
        for r in range(world):
            if need[r] > 0:
                reqs.append(dist.isend(torch.from_numpy(np.ascontiguousarray(gids[off[r]:off[r] + need[r]])), r))
        for r in range(world):
            buf = torch.zeros(int(send_cnt[r]), dtype=torch.int64)
            if send_cnt[r] > 0:
                dist.recv(buf, r)
            asked_parts.append(buf.numpy())
        for q in reqs:
            q.wait()
        asked = np.concatenate(asked_parts) if asked_parts else np.zeros(0, dtype=np.int64)
        sidx = np.zeros(max(len(asked), 1), dtype=np.int32)
        assert lib.b200_halo_send_index(n, world, rank, len(asked), np.ascontiguousarray(asked).ctypes.data_as(C.POINTER(C.c_int64)),
                                        sidx.ctypes.data_as(C.POINTER(C.c_int32))) == 0
        sidx = sidx[:len(asked)]
        # the multivector: partition-independent LCG stream (what b200_mv_random produces per rank)
        x_loc = oracle.random_rows(1, lo, hi, n, k)
        # halo exchange: pack rows, send per peer, receive into the ghost tail
        ghost = np.zeros((len(gids), k))
        reqs = []
        s_off = np.concatenate([[0], np.cumsum(send_cnt)])
        for r in range(world):
            if send_cnt[r] > 0:
                reqs.append(dist.isend(torch.from_numpy(np.ascontiguousarray(x_loc[sidx[s_off[r]:s_off[r + 1]]])), r))
        for r in range(world):
            if need[r] > 0:
                buf = torch.zeros((int(need[r]), k), dtype=torch.float64)
                dist.recv(buf, r)
                ghost[off[r]:off[r] + need[r]] = buf.numpy()
        for q in reqs:
            q.wait()
        y_loc = oracle.spmm(rp, loc, va, np.vstack([x_loc, ghost]))
        # serial oracle on the whole problem
        _, rpg, cig, vag = oracle.galeri_csr(name, *dims)
        xg, _ = oracle.random(1, n, k)
        yg = oracle.spmm(rpg, cig.astype(np.int32), vag, xg)
        assert np.array_equal(x_loc, xg[lo:hi]), "MvRandom stream is not partition independent"
        assert np.array_equal(y_loc, yg[lo:hi]), f"{name}: distributed SpMM differs from serial"
        # Gram block: local partial + allreduce == serial (to rounding), identical on every rank
        G = torch.from_numpy(np.ascontiguousarray(oracle.trans_mv(1.0, y_loc, x_loc)))
        dist.all_reduce(G)
        Gs = oracle.trans_mv(1.0, yg, xg)
        assert np.linalg.norm(G.numpy() - Gs) <= 1e-13 * np.linalg.norm(Gs)
        gl = [torch.zeros_like(G) for _ in range(world)]
        dist.all_gather(gl, G)
        assert all(torch.equal(g, gl[0]) for g in gl), "allreduced block differs between ranks"
    dist.barrier()
    if rank == 0:
        print("GLOO_WORKER_OK")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
