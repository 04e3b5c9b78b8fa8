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


def shard_of(lo: int, hi: int, rank: int, world: int) -> Tuple[int, int]:
    """Elements [first, first + count) of the flat range [lo, hi) that `rank` owns in the fused exchange + AdamW kernel
    (csrc/dp_exchange.cu, grappa_dp_shard): equal shares in units of 4 elements, so every shard stays 16-byte aligned."""
    if lo % 4 or (hi - lo) % 4:
        raise ValueError("the range must be a multiple of 4 elements")
    t4 = (hi - lo) // 4
    a, b = t4 * rank // world, t4 * (rank + 1) // world
    return lo + 4 * a, 4 * (b - a)


def update_stages(offsets, numel: int, n_blocks: int, ffn_split_blocks=(0,)):
    """Forward-ordered parameter stages of the staged exchange + update (train_step.TrainStep.exchange_and_update).

    `offsets`: parameter name -> first element in the flat buffers, in registration order (GNN first, then the writers).
    Returns (bounds, stage_of_mark, writer_stage, bucket_split):
      bounds         sorted stage boundaries, stage i = [bounds[i], bounds[i+1]);
      stage_of_mark  (block, part) -> stage whose parameters the forward reads first at that point: part 0 = the start of GNN
                     block i >= 1, part 1 = the self-interaction weights of a block listed in `ffn_split_blocks` (only when
                     they are the tail of the block's parameters, so that the cut leaves two contiguous ranges);
      writer_stage   stage of the four writers (None without writers); bucket_split = its first element."""
    names = list(offsets)
    writers = [n for n in names if not n.startswith("gnn.")]
    bucket_split = offsets[writers[0]] if writers else numel
    if not all(offsets[n] >= bucket_split for n in writers):
        raise ValueError("writer parameters must follow the GNN's")
    marks = {}
    for i in range(n_blocks):
        own = sorted((offsets[n], n) for n in names if n.startswith(f"gnn.att_blocks.{i}.") or n.startswith(f"gnn.blocks.{i}."))
        if not own:
            continue
        if i >= 1:
            marks[(i, 0)] = own[0][0]
        ffn = [o for o, n in own if ".self_interaction." in n]
        if i in ffn_split_blocks and ffn and all(".self_interaction." in n for o, n in own if o >= min(ffn)):
            marks[(i, 1)] = min(ffn)
    bounds = sorted(set([0] + list(marks.values()) + [bucket_split, numel]))
    stage_of_mark = {k: bounds.index(o) for k, o in marks.items()}
    writer_stage = bounds.index(bucket_split) if bucket_split in bounds[:-1] else None
    return bounds, stage_of_mark, writer_stage, bucket_split


def gather_shards_(tensors, ranges, group=None) -> None:
    """Completes every rank's copy of flat buffers whose ranges are sharded as `shard_of` says (the Adam moments under the
    fused exchange): each owner broadcasts its shard of every range.  In place; any backend."""
    dist = torch.distributed
    group = group if group is not None else dist.group.WORLD
    world = dist.get_world_size(group)
    for (lo, hi) in ranges:
        for r in range(world):
            first, count = shard_of(lo, hi, r, world)
            if count:
                src = dist.get_global_rank(group, r)
                for t in tensors:
                    dist.broadcast(t[first:first + count], src=src, group=group)


class SwitchMemory:
    """Flat buffers in NVSwitch symmetric memory: every rank allocates the same sizes, the driver maps all of them into
    one multicast address per buffer (a store to it lands in every rank's copy, a `multimem.ld_reduce` from it returns
    the sum over the copies, added in the switch).  torch.distributed's symmetric-memory allocator does the rendezvous;
    the kernels that use the addresses are ours (csrc/dp_exchange.cu).

    `available()` is False on one GPU, without NVSwitch multicast support, or when GRAPPA_DP_SWITCH=0: the training
    driver then exchanges gradients with NCCL all-reduces."""

    PAD_WORDS = 16384

    def __init__(self, device, group=None):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        self._symm = symm
        self.device = device
        self.group = group if group is not None else dist.group.WORLD
        self.rank = dist.get_rank(self.group)
        self.world = dist.get_world_size(self.group)
        self._handles = []
        pad = symm.empty(self.PAD_WORDS, dtype=torch.int32, device=device)
        pad.zero_()
        h = symm.rendezvous(pad, self.group)
        self._handles.append((pad, h))
        self.pad = pad
        self.pad_ptrs = [int(p) for p in h.buffer_ptrs]
        self.multicast = int(h.multicast_ptr) != 0
        torch.cuda.synchronize(device)
        dist.barrier(group=self.group)                  # every rank's flags are zero before anyone raises one

    @staticmethod
    def available(world: int) -> bool:
        import os
        if world <= 1 or os.environ.get("GRAPPA_DP_SWITCH", "1") == "0" or not torch.cuda.is_available():
            return False
        try:
            import importlib
            importlib.import_module("torch.distributed._symmetric_memory")
        except Exception:
            return False
        return True

    def zeros(self, numel: int):
        """(local tensor, multicast address) of a zero-filled fp32 buffer that exists on every rank."""
        t = self._symm.empty(numel, dtype=torch.float32, device=self.device)
        t.zero_()
        h = self._symm.rendezvous(t, self.group)
        self._handles.append((t, h))
        mc = int(h.multicast_ptr)
        if mc == 0:
            raise RuntimeError("no NVSwitch multicast mapping for this allocation")
        return t, mc
