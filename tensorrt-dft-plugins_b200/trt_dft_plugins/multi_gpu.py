"""Batch x channel sharding of the Rfft / Irfft path across the GPUs of one node.

Every image in the folded batch (dft_plugins.cpp:260-263) is independent, so multi-GPU execution is
"one process per GPU, each transforms a contiguous slice of the batch" with no collective in the data
path.  torch.distributed is used only for rendezvous, barriers, timing reductions and
the optional gather of results.  The helpers below hold that host-side logic so that it can be tested
with the gloo backend on CPU (tests/test_multi_gpu_gloo.py); the transform itself is passed in.
"""
from typing import Callable, Tuple


def shard_batch(batch: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous split of `batch` images over `world` ranks: returns (start, count) for `rank`.
    The first batch % world ranks get one extra image."""
    if world < 1 or not (0 <= rank < world) or batch < 0:
        raise ValueError("bad shard request")
    base, extra = divmod(batch, world)
    start = rank * base + min(rank, extra)
    return start, base + (1 if rank < extra else 0)


def max_over_ranks(value: float, device=None) -> float:
    """Max of a per-rank scalar (e.g. a CUDA-event time) over all ranks; identity without a process group."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sharded_apply(x_full_shape, signal_ndim: int, transform: Callable, make_input: Callable, gather: bool = False):
    """Runs `transform` on this rank's slice of a [batch..., signal dims] problem.

    make_input(start, count) -> this rank's input tensor with leading dim `count` (the folded batch);
    transform(x_slice)       -> the transformed slice.
    With gather=True the slices are collected on rank 0 (outside any timed region) and returned there.
    """
    import torch
    import torch.distributed as dist
    world = dist.get_world_size() if dist.is_initialized() else 1
    rank = dist.get_rank() if dist.is_initialized() else 0
    batch = 1
    for s in x_full_shape[:-signal_ndim]:
        batch *= int(s)
    start, count = shard_batch(batch, rank, world)
    y = transform(make_input(start, count))
    if not gather or world == 1:
        return y
    counts = [shard_batch(batch, r, world)[1] for r in range(world)]
    # gather needs equal sizes: pad every slice to the largest one and trim on rank 0
    cmax = max(counts)
    padded = torch.zeros((cmax,) + tuple(y.shape[1:]), dtype=y.dtype, device=y.device)
    padded[:count] = y
    pieces = [torch.empty_like(padded) for _ in range(world)] if rank == 0 else None
    dist.gather(padded, pieces, dst=0)
    return torch.cat([p[:c] for p, c in zip(pieces, counts)], dim=0) if rank == 0 else None
