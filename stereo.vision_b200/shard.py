"""Multi-GPU plumbing: one process per GPU, one independent sequence per rank.

Frames of a sequence are chained through prev_disp (loop.py:57,78) so a sequence never leaves its GPU;
different sequences share nothing but the read-only masks.  There is no collective on the path: the
only communication is the reduction of the per-rank timings after the timed region.
"""
import hashlib
import os

import numpy as np
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


def min_over_ranks(values, device="cpu"):
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    if dist.is_initialized():
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return t.tolist()


def sum_over_ranks(values, device="cpu"):
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    if dist.is_initialized():
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t.tolist()


def result_sha(road_mask, planes, counts) -> str:
    """sha256 over a sequence's bit-exact observables: the road masks (bit-packed), N / M / K / R per frame, the
    winning hypothesis and whether a plane was found.  A rank's value must equal the single-GPU value of the same
    sequence id; bench.py prints it per rank."""
    m = road_mask.cpu().numpy() if torch.is_tensor(road_mask) else np.asarray(road_mask)
    p = planes.cpu().numpy() if torch.is_tensor(planes) else np.asarray(planes)
    c = counts.cpu().numpy() if torch.is_tensor(counts) else np.asarray(counts)
    h = hashlib.sha256()
    h.update(np.packbits(m != 0).tobytes())
    h.update(np.ascontiguousarray(c, dtype=np.int32).tobytes())
    h.update(np.ascontiguousarray(p[:, 6], dtype=np.float64).astype(np.int32).tobytes())
    h.update((p[:, 7] != 0).astype(np.uint8).tobytes())
    return h.hexdigest()


def planes_sha(planes) -> str:
    """sha256 over the float64 bits of every frame's plane row (a, b, c, |abc|, cutoff, best error, ...): equal between
    two runs of the same kernels on the same sequence, whichever GPU ran it."""
    p = planes.cpu().numpy() if torch.is_tensor(planes) else np.asarray(planes)
    return hashlib.sha256(np.ascontiguousarray(p, dtype=np.float64).tobytes()).hexdigest()


def gather_strings(value: str, device="cpu"):
    """every rank's string on every rank (fixed 128-byte slots; hashes)"""
    rank, world, _ = rank_world()
    if not dist.is_initialized():
        return [value]
    buf = torch.zeros((128,), dtype=torch.uint8, device=device)
    b = value.encode()
    if len(b) > 128:
        raise ValueError("gather_strings carries at most 128 bytes per rank")
    buf[:len(b)] = torch.tensor(list(b), dtype=torch.uint8, device=device)
    out = [torch.zeros_like(buf) for _ in range(world)]
    dist.all_gather(out, buf)
    return [bytes(o.cpu().tolist()).rstrip(b"\0").decode() for o in out]


def finish():
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()
