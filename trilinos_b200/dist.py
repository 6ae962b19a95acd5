"""One process per GPU: context + NCCL communicator bootstrap and the row-block problem setup.

Replaces, for this path, Epetra_MpiComm construction + Galeri::CreateMap("Linear") +
Galeri::CreateCrsMatrix on that map (belos/epetra/example/BlockGmres/BlockGmresEpetraGaleriEx.cpp:103-152):
every rank owns the contiguous row block of Epetra_Map(n, 0, comm) (epetra/src/Epetra_BlockMap.cpp:83-90).
torch.distributed is used only as rendezvous plumbing (to ship the 128-byte NCCL unique id and for
barriers); the data path uses the library's own NCCL communicator.
"""
import os

from .core import Context, CrsOperator, MultiVec, galeri, linear_partition

DIMS3 = ("Laplace3D", "Brick3D")


def env_ranks():
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def init_from_env():
    """Returns (ctx, rank, world).  Under torchrun (WORLD_SIZE > 1) initialises torch.distributed
    (nccl when CUDA is there, gloo otherwise) and the library's NCCL communicator."""
    rank, world, local_rank = env_ranks()
    if world > 1:
        import torch
        import torch.distributed as dist
        if not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29500")
            if torch.cuda.is_available():
                torch.cuda.set_device(local_rank)
                dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local_rank))
            else:
                dist.init_process_group("gloo", rank=rank, world_size=world)
    ctx = Context(local_rank)
    if world > 1:
        import torch.distributed as dist
        box = [Context.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        ctx.comm_init(box[0], world, rank)
    return ctx, rank, world


def grid_size(name, dims):
    nx, ny, nz = dims
    return nx * ny * nz if name in DIMS3 else (nx if name in ("Tridiag", "Diag") else nx * ny)


def local_operator(ctx, name, dims, conv=1.0, diff=1.0e-5):
    """This rank's row block of the Galeri matrix as a CrsOperator.  Returns (op, n_global, row_begin, row_end)."""
    rank, world = ctx.rank
    n = grid_size(name, dims)
    lo, hi = linear_partition(n, world, rank)
    _, rp, ci, va = galeri(name, *dims, conv=conv, diff=diff, row_begin=lo, row_end=hi)
    return CrsOperator(ctx, n, lo, hi, rp, ci, va), n, lo, hi


def local_multivec(ctx, n_global, lo, hi, ncols, zero=True):
    return MultiVec(ctx, hi - lo, ncols, n_global=n_global, row_offset=lo, zero=zero)
