"""Multi-GPU partitioning of the hot path: frames / stereo pairs are independent units,
sharded across ranks (one process per GPU) with NO data-path collective.  torch.distributed is used
only for the launch plumbing: a barrier around timed regions and a max-over-ranks of the elapsed time.
"""
from __future__ import annotations


def shard_range(n_units: int, rank: int, world: int) -> range:
    """Contiguous, balanced shard of `n_units` (frames, or stereo pairs — a pair's two eyes must stay
    on one GPU because ComputeStereoMatches reads both pyramids)."""
    if not (0 <= rank < world) or n_units < 0:
        raise ValueError("bad shard arguments")
    base, rem = divmod(n_units, world)
    start = rank * base + min(rank, rem)
    return range(start, start + base + (1 if rank < rem else 0))


def _dist():
    import torch.distributed as dist
    return dist if dist.is_available() and dist.is_initialized() else None


def dist_reduce(value: float, op: str = "max") -> float:
    """max / sum of a scalar over ranks (identity when not distributed).  nccl needs a CUDA tensor."""
    dist = _dist()
    if dist is None:
        return float(value)
    import torch
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t = torch.tensor([float(value)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX if op == "max" else dist.ReduceOp.SUM)
    return float(t.item())


def barrier() -> None:
    dist = _dist()
    if dist is not None:
        dist.barrier()


def whole_job_throughput(units_this_rank: int, elapsed_ms_this_rank: float) -> float:
    """units all ranks processed / max elapsed over ranks (units per second)."""
    total = dist_reduce(units_this_rank, "sum")
    ms = dist_reduce(elapsed_ms_this_rank, "max")
    return total / (ms * 1e-3)
