"""
Data-parallel training-step driver.  Replaces, for benchmarking and as the minimal trainer, the
Lightning orchestration of training/lightning_model.py:228-254,368-370: forward (GrappaModel -> Energy), molecule-wise
loss, backward, gradient all-reduce over NCCL (one process per GPU), AdamW.

B200-first design choices
  - parameters live in ONE flat fp32 buffer and their gradients in another: the backward GEMMs write weight
    gradients straight into the flat gradient buffer (ops.GRAD_SINK), the all-reduce is a single NCCL call over
    83 MB, AdamW is a single kernel launch;
  - inputs arrive as two pinned host buffers (batching.PackedBatch) copied into static device buffers;
  - forward + loss + backward are captured once per batch shape in a CUDA graph and replayed; dropout masks stay
    fresh across replays through a device-side seed that the graph itself increments.
Molecules are independent units: rank r owns a contiguous slice of the global batch, no data-path collective exists
other than the gradient all-reduce.
"""
from typing import Dict, Optional

import torch

from . import _lib, ops
from .batching import DeviceBatch, PackedBatch
from .dp import all_reduce_sum_
from .energy import Energy
from .plan import get_plan

# parameters that the reference registers but never uses (perm_equiv_transformer.py:126): no gradient, not optimised
_UNUSED = ("grappa_transformer.norm.",)


class FlatParameters:
    def __init__(self, model: torch.nn.Module, device):
        self.named = [(n, p) for n, p in model.named_parameters() if p.requires_grad and not any(u in n for u in _UNUSED)]
        self.offsets: Dict[str, int] = {}
        total = 0
        for n, p in self.named:
            self.offsets[n] = total
            total += (p.numel() + 3) // 4 * 4            # 16-byte aligned views (vectorised GEMM loads)
        self.numel = total
        self.params = torch.zeros(total, device=device, dtype=torch.float32)
        self.grads = torch.zeros(total, device=device, dtype=torch.float32)
        self.exp_avg = torch.zeros(total, device=device, dtype=torch.float32)
        self.exp_avg_sq = torch.zeros(total, device=device, dtype=torch.float32)
        self.sink = {}
        for n, p in self.named:
            o = self.offsets[n]
            view = self.params[o:o + p.numel()].view_as(p)
            view.copy_(p.data)
            p.data = view
            self.sink[id(p)] = self.grads[o:o + p.numel()].view_as(p)

    def grad_of(self, name: str) -> torch.Tensor:
        p = dict(self.named)[name]
        return self.sink[id(p)]


class TrainStep:
    def __init__(self, model, device, lr: float = 1e-5, weight_decay: float = 0.0, betas=(0.9, 0.999), eps: float = 1e-8,
                 energy_weight: float = 1.0, gradient_weight: float = 0.8, proper_regularisation: float = 1.0,
                 improper_regularisation: float = 1.0, world_size: int = 1, use_cuda_graph: bool = True,
                 process_group=None, overlap_comm: Optional[bool] = None):
        self.device = ops.norm_device(device)
        self.model = model.to(self.device).train()
        self.energy = Energy(terms=["bond", "angle", "proper", "improper"], materialize_tuple_terms=False)
        self.flat = FlatParameters(self.model, self.device)
        self.lr, self.wd, self.betas, self.eps = lr, weight_decay, betas, eps
        self.weights = (energy_weight, gradient_weight, proper_regularisation, improper_regularisation)
        self.world_size = world_size
        self.pg = process_group
        self.use_cuda_graph = use_cuda_graph
        self.t = 0
        # Two gradient buckets in backward order: the four writers' parameters (second half of the flat
        # buffer) are complete once the backward pass reaches the atom embedding, so their all-reduce runs while the
        # GNN backward is still computing; the GNN bucket follows.  Needs the backward split at the embedding.
        self.overlap_comm = (world_size > 1) if overlap_comm is None else overlap_comm
        first_writer = [n for n, _ in self.flat.named if not n.startswith("gnn.")]
        self.bucket_split = self.flat.offsets[first_writer[0]] if first_writer else self.flat.numel
        assert all(self.flat.offsets[n] >= self.bucket_split for n in first_writer), "writer parameters must follow the GNN's"
        self.seed_dev = torch.zeros(1, dtype=torch.int64, device=self.device)
        self._static: Dict[tuple, dict] = {}
        # warm-up and capture both run on this side stream (the whole-network capture recipe of torch.cuda.graphs:
        # autograd's per-stream state must exist before capture begins)
        self._side = torch.cuda.Stream(device=self.device) if use_cuda_graph else None

    # ---- one forward + loss + backward on device-resident inputs -------------------------------------------------
    def _fwd_bwd(self, g):
        ops.GRAD_SINK = self.flat.sink
        ops.DropoutSeeds.dev = self.seed_dev
        try:
            g = self.energy(self.model(g))
            plan = get_plan(g)
            gd, n1 = g.nodes["g"].data, g.nodes["n1"].data
            loss = ops.MolwiseLossFn.apply(gd["energy"], n1["gradient"], g.nodes["n4"].data["k"], g.nodes["n4_improper"].data["k"],
                                           gd["energy_ref"], n1["gradient_ref"], gd.get("is_dummy"), plan, self.weights)
            loss.backward()
            ops.join_side_work(self.device)
            self.seed_dev.add_(1)
        finally:
            ops.GRAD_SINK = None
            ops.DropoutSeeds.dev = None
        return loss.detach()

    # ---- the same in two phases, split where the backward pass reaches the atom embedding ------------------------------
    def _phase_a(self, g):
        """GNN forward, writers + energy + loss forward, backward down to the embedding (all writer gradients done)."""
        ops.GRAD_SINK = self.flat.sink
        ops.DropoutSeeds.dev = self.seed_dev
        try:
            plan = get_plan(g)
            g = self.model.gnn(g, plan)
            h = g.nodes["n1"].data["h"]
            leaf = h.detach().requires_grad_(True)
            g.nodes["n1"].data["h"] = leaf
            g = self.energy(self.model.parameter_writer(g, plan))
            gd, n1 = g.nodes["g"].data, g.nodes["n1"].data
            loss = ops.MolwiseLossFn.apply(gd["energy"], n1["gradient"], g.nodes["n4"].data["k"], g.nodes["n4_improper"].data["k"],
                                           gd["energy_ref"], n1["gradient_ref"], gd.get("is_dummy"), plan, self.weights)
            loss.backward()
            # the backward pass ends on the four writer streams (nothing downstream consumes them in this phase):
            # join them explicitly, which is also what closes the forked branches under graph capture
            from . import model as _model
            if _model.PARALLEL_WRITERS:
                cur = torch.cuda.current_stream(self.device)
                for wst in _model._writer_streams(self.device):
                    cur.wait_stream(wst)
            ops.join_side_work(self.device)
        finally:
            ops.GRAD_SINK = None
            ops.DropoutSeeds.dev = None
        return loss.detach(), h, leaf

    def _phase_b(self, h, leaf):
        """GNN backward from the embedding gradient."""
        ops.GRAD_SINK = self.flat.sink
        ops.DropoutSeeds.dev = self.seed_dev
        try:
            h.backward(leaf.grad)
            ops.join_side_work(self.device)
            self.seed_dev.add_(1)
        finally:
            ops.GRAD_SINK = None
            ops.DropoutSeeds.dev = None

    def _run_split(self, st, g, between):
        """Phase A, `between()` (the first bucket's all-reduce), phase B; eager during warm-up, two graphs afterwards."""
        if not self.use_cuda_graph:
            loss, h, leaf = self._phase_a(g)
            between()
            self._phase_b(h, leaf)
            return loss
        if st.get("graph_a") is None:
            cur = torch.cuda.current_stream(self.device)
            if st["warm"] < 3:
                st["warm"] += 1
                self._side.wait_stream(cur)
                with torch.cuda.stream(self._side):
                    loss, h, leaf = self._phase_a(g)
                    self._phase_b(h, leaf)
                cur.wait_stream(self._side)
                between()
                return loss
            torch.cuda.synchronize(self.device)
            ga, gb = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
            with torch.cuda.graph(ga, stream=self._side):
                st["loss"], h, leaf = self._phase_a(g)
            with torch.cuda.graph(gb, stream=self._side, pool=ga.pool()):
                self._phase_b(h, leaf)
            st["graph_a"], st["graph_b"] = ga, gb
        st["graph_a"].replay()
        between()
        st["graph_b"].replay()
        return st["loss"]

    def _step_overlapped(self, st) -> torch.Tensor:
        g = st["batch"].graph
        works = []
        dist = torch.distributed
        multi = self.world_size > 1

        def first_bucket():
            if multi and self.bucket_split < self.flat.numel:
                works.append(dist.all_reduce(self.flat.grads[self.bucket_split:], op=dist.ReduceOp.SUM, group=self.pg, async_op=True))

        loss = self._run_split(st, g, first_bucket)
        if multi and self.bucket_split > 0:
            works.append(dist.all_reduce(self.flat.grads[:self.bucket_split], op=dist.ReduceOp.SUM, group=self.pg, async_op=True))
        for w in works:
            w.wait()
        self._grad_scale = 1.0 / self.world_size
        self.optimizer_step()
        return loss

    def _state_for(self, packed: PackedBatch):
        st = self._static.get(packed.signature)
        if st is None:
            st = {"batch": DeviceBatch(packed, self.device), "graph": None, "loss": None, "warm": 0}
            self._static[packed.signature] = st
        return st

    def compute_gradients(self, packed: PackedBatch, load: bool = True) -> torch.Tensor:
        """H2D (optional) + forward + loss + backward into the flat gradient buffer.  Returns the device loss."""
        st = self._state_for(packed)
        if load:
            st["batch"].load(packed)
        g = st["batch"].graph
        if not self.use_cuda_graph:
            return self._fwd_bwd(g)
        if st["graph"] is None:
            cur = torch.cuda.current_stream(self.device)
            if st["warm"] < 3:                      # eager warm-up: lazy host constants, smem attributes, allocator
                st["warm"] += 1
                self._side.wait_stream(cur)
                with torch.cuda.stream(self._side):
                    loss = self._fwd_bwd(g)
                cur.wait_stream(self._side)
                return loss
            torch.cuda.synchronize(self.device)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=self._side):
                st["loss"] = self._fwd_bwd(g)
            st["graph"] = graph
        st["graph"].replay()
        return st["loss"]

    def all_reduce_gradients(self):
        self._grad_scale = all_reduce_sum_(self.flat.grads, self.world_size, self.pg)

    def optimizer_step(self):
        self.t += 1
        f = self.flat
        _lib.check(_lib.lib().grappa_adamw(f.numel, ops._p(f.params), ops._p(f.grads), ops._p(f.exp_avg), ops._p(f.exp_avg_sq),
                                           self.lr, self.betas[0], self.betas[1], self.eps, self.wd, self.t,
                                           getattr(self, "_grad_scale", 1.0 / self.world_size), ops._stream()))

    def step_from_store(self, store, mol_ids, conf_strategy="max", generator=None) -> torch.Tensor:
        """One optimisation step on a batch collated ON THE DEVICE from a resident `store.MoleculeStore`: the host only
        uploads the segment tables (tens of kB), not the batch."""
        spec = store.prepare(mol_ids, conf_strategy, generator)
        st = self._state_for(spec)
        store.collate_into(spec, st["batch"])
        if self.overlap_comm:
            return self._step_overlapped(st)
        loss = self.compute_gradients(spec, load=False)
        self.all_reduce_gradients()
        self.optimizer_step()
        return loss

    def step(self, packed: PackedBatch, load: bool = True) -> torch.Tensor:
        if self.overlap_comm:
            st = self._state_for(packed)
            if load:
                st["batch"].load(packed)
            return self._step_overlapped(st)
        loss = self.compute_gradients(packed, load=load)
        self.all_reduce_gradients()
        self.optimizer_step()
        return loss
