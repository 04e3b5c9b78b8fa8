"""
Data-parallel plumbing that does not depend on the device kernels (so it is testable with the gloo backend on CPU):
which molecules a rank owns, and the gradient exchange over a flat buffer.

The path shards by molecules (independent units: no edges across molecules after batching, utils/dgl_utils.py:11-63;
the loss is a plain mean over molecules, training/loss.py:172).  With equal molecule counts per rank,
  mean over ranks of (local mean loss)  ==  global mean loss,
so the exchange is ONE all-reduce (sum) of the flat gradient buffer followed by a 1/world scale, which the AdamW
kernel applies for free (`grad_scale`).  Unequal counts need the weights n_r / n (not used by bench.py).
"""
from typing import List, Sequence, Tuple

import torch


def rank_slice(n_global: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous slice [lo, hi) of the global batch owned by `rank`; sizes differ by at most one."""
    base, rem = divmod(n_global, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def rank_seeds(rank: int, per_rank: int) -> List[int]:
    """Synthetic-data seeds of a rank (SURVEY.md §8d cfg 5: rank r uses seeds 32r .. 32r+31)."""
    return list(range(rank * per_rank, (rank + 1) * per_rank))


def flat_views(tensors: Sequence[torch.Tensor], align: int = 4):
    """One flat fp32 buffer with a 16-byte aligned view per tensor. Returns (flat, views, offsets)."""
    offsets, total = [], 0
    for t in tensors:
        offsets.append(total)
        total += (t.numel() + align - 1) // align * align
    device = tensors[0].device if tensors else "cpu"
    flat = torch.zeros(total, dtype=torch.float32, device=device)
    views = [flat[o:o + t.numel()].view_as(t) for o, t in zip(offsets, tensors)]
    return flat, views, offsets


def all_reduce_sum_(flat: torch.Tensor, world: int, group=None) -> float:
    """In-place sum over ranks; returns the scale (1/world) the optimizer must apply to obtain the global-mean gradient."""
    if world > 1:
        torch.distributed.all_reduce(flat, op=torch.distributed.ReduceOp.SUM, group=group)
    return 1.0 / world
