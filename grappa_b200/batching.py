"""
Packed batches: the host side of the training / inference boundary.

The reference ships a pickled DGL graph per step from DataLoader workers and moves it with `g.to(device)`, one copy
per tensor (training/lightning_model.py:228-236, grappa.py:67).  Here the collate step packs everything the hot path
reads into TWO pinned host buffers (one fp32, one int32: the batch plan plus the int64 idxs narrowed to int32), so a
step's input transfer is two `cudaMemcpyAsync` calls into static device buffers whose views form the device graph.
Static buffers are what lets the whole step be replayed as a CUDA graph.
"""
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from .graph import LEVELS, HeteroGraph
from .plan import BatchPlan, build_plan

_FLOAT_FIELDS = (("n1", "xyz"), ("n1", "gradient_ref"), ("g", "energy_ref"), ("g", "is_dummy"))


class PackedBatch:
    """Host-resident packed batch.  `fbuf` fp32, `plan.host` int32 (both pinned when requested)."""

    def __init__(self, g: HeteroGraph, feature_names: Optional[List[str]] = None, pin: bool = True,
                 plan: Optional[BatchPlan] = None):
        self.num_nodes = dict(g._num_nodes)
        self.batch_num_nodes = {k: v.clone() for k, v in g._batch_num_nodes.items()}
        self.n_edges = g.num_edges()
        n1 = g.nodes["n1"].data
        names = feature_names if feature_names is not None else [
            k for k in n1.keys() if k not in ("xyz", "gradient_ref", "ids", "h", "gradient") and torch.is_floating_point(n1[k])]
        fields: List[Tuple[str, str, torch.Tensor]] = [("n1", k, n1[k]) for k in names]
        for nt, k in _FLOAT_FIELDS:
            if k in g.nodes[nt].data:
                fields.append((nt, k, g.nodes[nt].data[k]))
        self.layout: Dict[Tuple[str, str], Tuple[int, Tuple[int, ...]]] = {}
        total = 0
        for nt, k, t in fields:
            self.layout[(nt, k)] = (total, tuple(t.shape))
            total += (t.numel() + 3) // 4 * 4
        fbuf = torch.empty(max(total, 4), dtype=torch.float32)
        for nt, k, t in fields:
            off, shape = self.layout[(nt, k)]
            fbuf[off:off + t.numel()] = t.reshape(-1).float()
        self.plan = plan if plan is not None else build_plan(g)
        if pin and torch.cuda.is_available():
            fbuf = fbuf.pin_memory()
            self.plan = self.plan.pin_memory()
        self.fbuf = fbuf
        self.edges = g._edges      # kept for containers that want the raw edge list (not needed by the kernels)

    @property
    def signature(self):
        """Shapes that must be equal for two batches to share static device buffers / a captured graph."""
        return (tuple(sorted(self.num_nodes.items())), tuple(sorted((k, v) for k, v in self.layout.items())),
                tuple(sorted(self.plan.offsets.items())), tuple(sorted(self.plan.energy_sizes.items())))

    @property
    def h2d_bytes(self) -> int:
        return int(self.fbuf.numel()) * 4 + self.plan.h2d_bytes

    @property
    def n_float(self) -> int:
        return int(self.fbuf.numel())

    @property
    def n_int(self) -> int:
        return int(self.plan.host.numel())


class DeviceBatch:
    """Static device buffers + a HeteroGraph whose tensors are views into them.  `packed` is a `PackedBatch` (filled by
    two host->device copies) or a `store.CollatedSpec` (filled on the device by `MoleculeStore.collate_into`)."""

    def __init__(self, packed, device):
        self.device = torch.device(device)
        self.signature = packed.signature
        self.fbuf = torch.empty(packed.n_float, dtype=torch.float32, device=self.device)
        self.ibuf = torch.empty(packed.n_int, dtype=torch.int32, device=self.device)
        plan = BatchPlan()
        plan.__dict__.update(packed.plan.__dict__)
        plan.dev = self.ibuf
        self.plan = plan
        g = HeteroGraph(packed.num_nodes, packed.edges, packed.batch_num_nodes)
        for (nt, k), (off, shape) in packed.layout.items():
            n = int(np.prod(shape)) if len(shape) else 1
            g.nodes[nt].data[k] = self.fbuf[off:off + n].view(shape)
        for lvl in LEVELS:
            if lvl in g.ntypes:
                # int32 device view of the tuple indices (the int64 originals stay on the host)
                g.nodes[lvl].data["idxs"] = plan.view("idx_" + lvl).view(-1, {"n2": 2, "n3": 3}.get(lvl, 4))
        g._plan_cache["plan"] = plan
        self.graph = g
        if isinstance(packed, PackedBatch):
            self.load(packed)

    def load(self, packed: PackedBatch, non_blocking: bool = True):
        """Two host->device copies on the current stream."""
        if packed.signature != self.signature:
            raise ValueError("batch shapes differ from the static device buffers")
        self.fbuf.copy_(packed.fbuf, non_blocking=non_blocking)
        self.ibuf.copy_(packed.plan.host, non_blocking=non_blocking)
        self.refresh_host_fields(packed)
        return self.graph

    def refresh_host_fields(self, packed):
        """The signature fixes totals and offsets, not the per-molecule split: host-side readers of
        `g.batch_num_nodes()` / `plan.mol_atom_counts` (unbatching, per-molecule metrics) must see THIS batch's."""
        self.graph._batch_num_nodes = {k: v.clone() for k, v in packed.batch_num_nodes.items()}
        self.plan.mol_atom_counts = packed.plan.mol_atom_counts
