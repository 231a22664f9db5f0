"""Range arithmetic for running the path on several GPUs.

Every shot and every fault path is independent (the reference re-zeroes the frame per
``propagate`` call, errorpropagator.cpp:262-263), so the path shards with no data-path
collective: rank r owns one contiguous range of shot ids / path indices, and only the
tallies (per-clbit marginals, counts, fault histogram: ``uint64[nc + 66]``) are summed, with
one all-reduce.  One process per GPU; ``torch.distributed`` (NCCL on GPUs, gloo in the CPU
tests) is the plumbing.
"""

from __future__ import annotations

from typing import Tuple

LANE = 32  # paths / shots per lane word


def shard_range(total: int, rank: int, world: int, align: int = LANE) -> Tuple[int, int]:
    """[lo, hi) of ``rank`` when [0, total) is cut into ``world`` contiguous pieces whose
    boundaries are multiples of ``align`` (so that every rank computes whole lane words and
    the concatenation of the ranks' packed outputs is the single-GPU output).  Pieces differ
    by at most ``align`` elements; trailing ranks may be empty when total < world * align."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("rank must be in [0, world)")
    if total < 0 or align < 1:
        raise ValueError("total must be >= 0 and align >= 1")
    units = -(-total // align)  # lane words
    lo_u = units * rank // world
    hi_u = units * (rank + 1) // world
    return min(lo_u * align, total), min(hi_u * align, total)


def shard_shots(shot_lo: int, n_shots: int, rank: int, world: int) -> Tuple[int, int]:
    """(first shot id, number of shots) of ``rank`` for the global shot range
    [shot_lo, shot_lo + n_shots); shot_lo must be a multiple of 32.  The sampler's Philox
    counter is the GLOBAL 32-shot word index, so the union of the ranks' results is bit for
    bit the single-GPU result."""
    if shot_lo % LANE:
        raise ValueError("shot_lo must be a multiple of 32")
    lo, hi = shard_range(n_shots, rank, world)
    return shot_lo + lo, hi - lo


def allreduce_tally(tally, group=None):
    """Sum a tally tensor (int64, QEC_TALLY_LEN(nc)) over the ranks in place and return it.
    No-op without an initialised process group (single GPU)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(tally, op=dist.ReduceOp.SUM, group=group)
    return tally
