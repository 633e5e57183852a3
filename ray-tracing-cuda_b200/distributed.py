"""Multi-GPU host logic: one process per GPU, samples partitioned across ranks, ONE reduce.

Mirrors DistributedMain (utils.cu:181-242): every rank holds the whole scene, renders
GetWorkload(rank, world, spp) samples per pixel (utils.cu:111-113) and the per-rank radiance sums
are added at the root.  Differences by design:
  * rank r renders the GLOBAL sample indices [begin_r, begin_r + count_r) — the Philox key is
    (pixel, global sample, bounce), so N ranks compute the same estimator as one (the reference
    seeds every rank identically and renders duplicates);
  * the reduce runs on the device over NVLink, not as a D2H copy + host MPI_Reduce (utils.cu:124,232-238):
    one kernel per rank over CUDA-IPC peer frames that sums, resolves and gathers (open_peer_frame +
    rt_reduce_resolve_peers, the default), or one ncclReduce of the exact sums (bind_nccl + rt_reduce_fixed).
torch.distributed is plumbing only: rendezvous, the handle / ncclUniqueId exchange, barriers.
"""
import os

import numpy as np

from . import core


def env_rank():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def partition(spp, rank, world_size):
    """(count, begin) of rank's global sample range — GetWorkload through the C-ABI."""
    return core.get_workload(rank, world_size, spp)


def init_process_group(backend=None):
    import torch
    import torch.distributed as dist

    rank, world, local = env_rank()
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        kw = {}
        if backend == "nccl":
            torch.cuda.set_device(local)
            kw["device_id"] = torch.device("cuda", local)
        dist.init_process_group(backend=backend, rank=rank, world_size=world, **kw)
    return rank, world, local


def broadcast_bytes(payload, n, src=0):
    """Ship `n` bytes from rank `src` to every rank over the default process group."""
    import torch
    import torch.distributed as dist

    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t = torch.zeros(n, dtype=torch.uint8, device=dev)
    if dist.get_rank() == src:
        t.copy_(torch.frombuffer(bytearray(payload), dtype=torch.uint8))
    dist.broadcast(t, src)
    return bytes(t.cpu().numpy().tobytes())


def bind_nccl(ctx):
    """Create the context's own NCCL communicator (rank 0 makes the id, everyone joins)."""
    import torch.distributed as dist

    rank, world = dist.get_rank(), dist.get_world_size()
    uid = core.nccl_unique_id() if rank == 0 else b"\0" * 128
    uid = broadcast_bytes(uid, 128, 0)
    ctx.nccl_init(uid, world, rank)


def open_peer_frame(ctx, height, width):
    """Every rank allocates its frame, the 64-byte IPC handles travel over the default process group, every rank maps
    the others (rt_peer.cu).  Returns the (sums, accum, rgb) device addresses of this rank's frame."""
    import torch.distributed as dist

    rank, world = (dist.get_rank(), dist.get_world_size()) if dist.is_available() and dist.is_initialized() else (0, 1)
    mine = ctx.peer_frame_create(height, width)
    handles = [broadcast_bytes(mine if rank == r else b"\0" * 64, 64, r) for r in range(world)] if world > 1 else [mine]
    ctx.peer_frame_open(world, rank, handles)
    return ctx.peer_frame_buffers()


def reduce_on_cpu(local_accum, root=0):
    """The same reduce for CPU-side tests of the partition logic (gloo): SUM to root."""
    import torch
    import torch.distributed as dist

    t = torch.from_numpy(np.ascontiguousarray(local_accum, np.float32))
    dist.reduce(t, root, op=dist.ReduceOp.SUM)
    return t.numpy()
