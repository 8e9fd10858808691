"""Multi-GPU plumbing: one process per GPU, one independent sequence per rank.

Frames of a sequence are chained through prev_disp (loop.py:57,78) so a sequence never leaves its GPU;
different sequences share nothing but the read-only masks.  There is no collective on the path: the
only communication is the reduction of the per-rank timings after the timed region.
"""
import os

import torch
import torch.distributed as dist


def rank_world():
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def init(backend: str, device=None):
    """Join the job described by RANK/WORLD_SIZE/MASTER_* (torchrun).  No-op for a single process."""
    rank, world, _ = rank_world()
    if world > 1 and not dist.is_initialized():
        kw = {"device_id": device} if (backend == "nccl" and device is not None) else {}
        dist.init_process_group(backend, **kw)
    return rank, world


def sequence_of_rank(rank: int, base: int = 0) -> int:
    """Sequence id processed by `rank` (weak scaling: every rank gets its own full sequence)."""
    return base + rank


def barrier():
    if dist.is_initialized():
        dist.barrier()


def max_over_ranks(values, device="cpu"):
    """Element-wise MAX over ranks of a short list of floats (device-timed milliseconds)."""
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    if dist.is_initialized():
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.tolist()


def sum_over_ranks(values, device="cpu"):
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    if dist.is_initialized():
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t.tolist()


def finish():
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()
