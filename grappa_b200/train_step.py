"""
Data-parallel training-step driver.  Replaces, for benchmarking and as the minimal trainer, the
Lightning orchestration of training/lightning_model.py:228-254,368-370: forward (GrappaModel -> Energy), molecule-wise
loss, backward, gradient exchange between the ranks (one process per GPU), AdamW.

B200-first design choices
  - parameters live in ONE flat fp32 buffer and their gradients in another: the backward GEMMs write weight
    gradients straight into the flat gradient buffer (ops.GRAD_SINK), AdamW is one launch per parameter stage;
  - for world_size > 1 the two flat buffers live in NVSwitch symmetric memory and the gradient exchange is FUSED with
    AdamW (csrc/dp_exchange.cu: in-switch sum of the owner's shard, update, multicast store of the new parameters),
    staged in forward order so that all stages but the first run under the next step's forward; NCCL all-reduces in
    the same staging where multicast memory is not available;
  - inputs arrive as two pinned host buffers (batching.PackedBatch) copied into static device buffers;
  - forward + loss + backward are captured once per batch shape in a CUDA graph and replayed; dropout masks stay
    fresh across replays through a device-side seed that the graph itself increments.
Molecules are independent units: rank r owns a contiguous slice of the global batch, no data-path collective exists
other than the gradient exchange.
"""
import ctypes as _ctypes
import os as _os
from typing import Dict, Optional

import torch

from . import _lib, ops
from .batching import DeviceBatch, PackedBatch
from .dp import SwitchMemory, all_reduce_sum_, gather_shards_, shard_of, update_stages
from .energy import Energy
from .plan import get_plan

# parameters that the reference registers but never uses (perm_equiv_transformer.py:126): no gradient, not optimised
_UNUSED = ("grappa_transformer.norm.",)


class _BucketBoundary(torch.autograd.Function):
    """Identity on the atom embedding between the GNN and the four writers.  Its backward runs once all writers'
    contributions to dL/dh have been accumulated, i.e. after every writer gradient kernel has been issued: the point
    where the writers' gradient bucket can go on the wire."""

    @staticmethod
    def forward(ctx, h, step):
        ctx.step = step
        return h.view_as(h)

    @staticmethod
    def backward(ctx, grad):
        ctx.step._reduce_writer_bucket()
        return grad, None


class _LossHandle:
    def __init__(self, buf, ev):
        self._buf, self._ev = buf, ev

    def value(self) -> float:
        self._ev.synchronize()
        return float(self._buf[0])


class FlatParameters:
    def __init__(self, model: torch.nn.Module, device, switch=None):
        self.named = [(n, p) for n, p in model.named_parameters() if p.requires_grad and not any(u in n for u in _UNUSED)]
        self.offsets: Dict[str, int] = {}
        total = 0
        for n, p in self.named:
            self.offsets[n] = total
            total += (p.numel() + 3) // 4 * 4            # 16-byte aligned views (vectorised GEMM loads)
        self.numel = total
        self.params_mc = self.grads_mc = 0
        if switch is not None:          # NVSwitch symmetric memory: the fused exchange + AdamW kernel addresses all ranks' copies
            self.params, self.params_mc = switch.zeros(total)
            self.grads, self.grads_mc = switch.zeros(total)
        else:
            self.params = torch.zeros(total, device=device, dtype=torch.float32)
            self.grads = torch.zeros(total, device=device, dtype=torch.float32)
        self.exp_avg = torch.zeros(total, device=device, dtype=torch.float32)
        self.exp_avg_sq = torch.zeros(total, device=device, dtype=torch.float32)
        self.params_lo = torch.zeros(total, device=device, dtype=torch.float32)     # lo planes of the pre-split weights
        self.lo_of = {}
        self.sink = {}
        for n, p in self.named:
            o = self.offsets[n]
            view = self.params[o:o + p.numel()].view_as(p)
            view.copy_(p.data)
            p.data = view
            self.sink[id(p)] = self.grads[o:o + p.numel()].view_as(p)
            self.lo_of[id(p)] = self.params_lo[o:o + p.numel()].view_as(p)

    def grad_of(self, name: str) -> torch.Tensor:
        p = dict(self.named)[name]
        return self.sink[id(p)]


class TrainStep:
    def __init__(self, model, device, lr: float = 1e-5, weight_decay: float = 0.0, betas=(0.9, 0.999), eps: float = 1e-8,
                 energy_weight: float = 1.0, gradient_weight: float = 0.8, proper_regularisation: float = 1.0,
                 improper_regularisation: float = 1.0, world_size: int = 1, use_cuda_graph: bool = True,
                 process_group=None, overlap_comm: Optional[bool] = None, split_update: Optional[bool] = None,
                 max_cached_shapes: int = 8, warmup_steps: int = 0, ffn_split_blocks=(0,),
                 switch_exchange: Optional[bool] = None):
        """`warmup_steps`: linear learning-rate warm-up from 0 to `lr` over that many steps, as
        training/lightning_model.py:136-161 (`_get_lr`; the reference's config uses 500; call `restart_warmup()` where
        the reference re-arms it at a restart epoch, :199-201).  0 (default) = constant lr, which is what bench.py
        times.  The epoch-level `lr_decay` of the reference's scheduler lives above this driver: assign `self.lr`."""
        self.device = ops.norm_device(device)
        self.warmup_steps = int(warmup_steps)
        self._warmup_step = 0 if self.warmup_steps > 0 else None
        self.model = model.to(self.device).train()
        self.energy = Energy(terms=["bond", "angle", "proper", "improper"], materialize_tuple_terms=False)
        # Gradient exchange for world_size > 1: on an NVSwitch domain with multicast memory the flat parameter and gradient
        # buffers live in symmetric memory and every parameter stage is exchanged AND updated by one kernel of ours
        # (grappa_dp_exchange_adamw: in-switch gradient sum -> AdamW on this rank's shard -> multicast store of the new
        # parameters; optimiser state sharded over the ranks, see gather_optimizer_state).  Otherwise: NCCL all-reduce per
        # stage + AdamW on every rank.
        self._switch = None
        self.pg = process_group
        if world_size > 1 and (switch_exchange is None or switch_exchange) and SwitchMemory.available(world_size):
            try:
                sw = SwitchMemory(self.device, process_group)
                self._switch = sw if sw.multicast else None
            except Exception as e:          # no multicast support on this box / driver: the NCCL path is complete on its own
                if switch_exchange:
                    raise
                import warnings
                warnings.warn(f"NVSwitch symmetric memory is not available ({str(e)[:200]}); gradients go through NCCL all-reduce")
        self._slot_of = {}
        # CTAs of the exchange kernel (it runs under the next forward: enough loads in flight for the links, few enough SMs
        # taken from the forward).  Measured on the cfg-2 step: 2 GPUs 16/32/64 CTAs -> 3.32/3.26/3.30 ms, 8 GPUs
        # 16/32/64 -> 3.25/3.30/3.31 ms (shards shrink with the world size).
        self._dp_ctas = int(_os.environ.get("GRAPPA_DP_CTAS", str(max(8, min(32, 128 // max(world_size, 1))))))
        self._dp_spin_seconds = float(_os.environ.get("GRAPPA_DP_SPIN_SECONDS", "120"))
        self.flat = FlatParameters(self.model, self.device, switch=self._switch)
        if ops.USE_WEIGHT_LO:                                   # lo planes of the initial weights (refreshed after every AdamW)
            _lib.check(_lib.lib().grappa_split_lo(self.flat.numel, ops._p(self.flat.params), ops._p(self.flat.params_lo),
                                                  ops._stream()))
        # bf16 planes of every weight matrix for the bf16x3 GEMMs (ops.MATH_MODE 2): one flat buffer, refreshed by ONE launch
        # after every AdamW (grappa_split_bf16 over a device-resident table of the matrices in the updated range)
        self._plane_tables = {}
        self._planes = None
        if ops.MATH_MODE == 2:
            self._init_planes()
        self.lr, self.wd, self.betas, self.eps = lr, weight_decay, betas, eps
        self.weights = (energy_weight, gradient_weight, proper_regularisation, improper_regularisation)
        self.world_size = world_size
        self.pg = process_group
        self.use_cuda_graph = use_cuda_graph
        self.t = 0
        # Two gradient buckets in backward order: the four writers' parameters (second half of the flat
        # buffer) are complete once the backward pass reaches the atom embedding, so their all-reduce is issued from
        # an autograd boundary node at that point, on a communication stream, and runs while the GNN backward is still
        # computing; the GNN bucket follows after backward.  Under CUDA-graph capture the NCCL launches become nodes of
        # the step's graph (a parallel branch).
        # STATUS (round 1): the stream choreography is verified on one GPU with a stand-in exchange
        # (tests/test_gpu_model.py::test_bucketed_gradient_exchange_choreography), but with real NCCL collectives the
        # 2-GPU run hung (twice, issued from the autograd thread and from the calling thread), so it is OFF by default and
        # refused for world_size > 1 (NCCL collectives captured into a CUDA graph hang on this stack: scratch/nccl_graph_probe.py).
        # Round 2 replaced it by the forward-ordered staged exchange below (split_update), fused with AdamW over NVSwitch memory.
        self.overlap_comm = bool(overlap_comm)
        if self.overlap_comm and world_size > 1:
            raise NotImplementedError("overlap_comm with NCCL hung on 2 x B200 in round 1; use the default single all-reduce")
        self._comm_stream = torch.cuda.Stream(device=self.device)
        self._copy_stream = torch.cuda.Stream(device=self.device)        # input prefetch (H2D of the next step's batch)
        # Split (staged) update: the first parameter stage is exchanged + updated on the compute stream, the later ones on an
        # update stream UNDER the next step's forward, which reads them later (stage list below; external-event wait nodes
        # in the step's graph).  On one GPU there is only an AdamW to hide and the wait nodes cost as much (3.237 vs 3.224 ms),
        # hence the default: on for world_size > 1.
        self.split_update = (world_size > 1) if split_update is None else bool(split_update)
        self._upd_stream = torch.cuda.Stream(device=self.device, priority=-1)
        self._writers_ready = torch.cuda.Event(external=True)
        self._writers_ready.record(torch.cuda.current_stream(self.device))
        self._stage_ready = []          # filled below once the stage boundaries are known
        self._loss_host = [torch.zeros(1).pin_memory() if torch.cuda.is_available() else torch.zeros(1) for _ in range(4)]
        self._loss_slot = 0
        self._works = []
        self._reduce_fn = None            # test hook: replaces the NCCL all-reduce of one bucket
        # flat-buffer range of every writer's parameters: a writer that is skipped in forward (level absent or empty in
        # this batch) never overwrites its gradient slices, which would otherwise still hold the LAST step's all-reduced
        # sums when the exchange and AdamW run over the whole buffer -> cleared at the start of such a step
        self._writer_ranges = {}
        for lvl, prefix in (("n2", "parameter_writer.bond_writer."), ("n3", "parameter_writer.angle_writer."),
                            ("n4", "parameter_writer.proper_writer."), ("n4_improper", "parameter_writer.improper_writer.")):
            offs = [(self.flat.offsets[n], p.numel()) for n, p in self.flat.named if n.startswith(prefix)]
            if offs:
                self._writer_ranges[lvl] = (min(o for o, _ in offs), max(o + (k + 3) // 4 * 4 for o, k in offs))
        # Staged update (world_size > 1): the parameters are exchanged + updated in FORWARD order in stages --
        # [pre_dense + GNN block 0 | block 1 | block 2 | block 3 + post_dense | the four writers].  Stage 0 runs on the compute
        # stream; the later ones (exchange + AdamW + plane refresh) on the update stream UNDER the next step's forward,
        # which waits for stage i only right before it reads that stage's parameters (external-event wait nodes in the
        # step graph).  Exposed communication = the first stage's exchange.
        # The block listed in `ffn_split_blocks` (default: block 0) is cut once more in front of its self-interaction
        # weights (4/5 of a block's parameters), so the one exchange nothing can hide -- the first stage's, which the next
        # forward needs at once -- carries pre_dense + the attention half of block 0 only (2.3 MB instead of 10.7 MB).
        n_blocks = len(getattr(self.model.gnn, "blocks", []))
        self._stage_bounds, self._stage_of_mark, self._writer_stage, self.bucket_split = update_stages(
            self.flat.offsets, self.flat.numel, n_blocks, tuple(ffn_split_blocks))
        for _ in range(len(self._stage_bounds) - 1):
            ev = torch.cuda.Event(external=True)
            ev.record(torch.cuda.current_stream(self.device))
            self._stage_ready.append(ev)
        self.seed_dev = torch.zeros(1, dtype=torch.int64, device=self.device)
        # static device buffers + captured graph per batch SHAPE (signature), least recently used first.  Real datasets
        # have many shapes: only `max_cached_shapes` stay resident (each holds its activations' memory pool).
        self._static: Dict[tuple, dict] = {}
        self.max_cached_shapes = max_cached_shapes
        # warm-up and capture both run on this side stream (the whole-network capture recipe of torch.cuda.graphs:
        # autograd's per-stream state must exist before capture begins)
        # priority -1: the activation-gradient chain goes ahead of the parameter-gradient side streams (priority 0)
        self._side = torch.cuda.Stream(device=self.device, priority=-1) if use_cuda_graph else None

    # ---- one forward + loss + backward on device-resident inputs -------------------------------------------------
    def _fwd_bwd(self, g):
        ops.GRAD_SINK = self.flat.sink
        # (bf16 mode: the flat tf32 lo planes are not maintained; the rare fallback GEMM makes its lo plane on demand)
        ops.WEIGHT_LO = self.flat.lo_of if (ops.USE_WEIGHT_LO and self._planes is None) else None
        ops.WEIGHT_PLANES = self.flat.planes_of if self._planes is not None else None
        ops.DropoutSeeds.dev = self.seed_dev
        try:
            plan = get_plan(g)

            def clear_skipped_writers():
                for lvl, (lo, hi) in self._writer_ranges.items():
                    if lvl not in g.ntypes or plan.T[lvl] == 0:
                        self.flat.grads[lo:hi].zero_()          # (a memset node of the captured step)
            if not self.split_update:
                clear_skipped_writers()
            if self.overlap_comm:
                plan = get_plan(g)
                g = self.model.gnn(g, plan)
                g.nodes["n1"].data["h"] = _BucketBoundary.apply(g.nodes["n1"].data["h"], self)
                g = self.energy(self.model.parameter_writer(g, plan))
            elif self.split_update:
                plan = get_plan(g)
                cur = torch.cuda.current_stream(self.device)
                self.model.gnn.stage_hook = lambda i, part: (self._stage_ready[self._stage_of_mark[(i, part)]].wait(cur)
                                                             if (i, part) in self._stage_of_mark else None)
                try:
                    g = self.model.gnn(g, plan)
                finally:
                    self.model.gnn.stage_hook = None
                if self._writer_stage is not None:
                    self._stage_ready[self._writer_stage].wait(cur)                  # writers' parameters of the last update
                clear_skipped_writers()      # only now: the deferred AdamW of the last step has finished reading the gradients
                g = self.energy(self.model.parameter_writer(g, plan))
            else:
                g = self.energy(self.model(g))
            plan = get_plan(g)
            gd, n1 = g.nodes["g"].data, g.nodes["n1"].data
            loss = ops.MolwiseLossFn.apply(gd["energy"], n1["gradient"], g.nodes["n4"].data["k"], g.nodes["n4_improper"].data["k"],
                                           gd["energy_ref"], n1["gradient_ref"], gd.get("is_dummy"), plan, self.weights)
            loss.backward()
            ops.join_side_work(self.device)
            if self.overlap_comm:
                # Both collectives are ISSUED here, by the calling thread (NCCL calls from the autograd worker thread
                # hung); what makes the first one overlap the GNN backward is its dependency: the communication
                # stream only waits for the writers' work as it stood when the boundary node ran.
                with torch.cuda.stream(self._comm_stream):
                    self._reduce_bucket(self.flat.grads[self.bucket_split:])      # writers' bucket
                self._reduce_bucket(self.flat.grads[:self.bucket_split])          # GNN bucket, after the whole backward
                cur = torch.cuda.current_stream(self.device)
                for w in self._works:
                    if w is not None:
                        w.wait()
                self._works.clear()
                cur.wait_stream(self._comm_stream)
            self.seed_dev.add_(1)
        finally:
            ops.GRAD_SINK = None
            ops.WEIGHT_LO = None
            ops.WEIGHT_PLANES = None
            ops.DropoutSeeds.dev = None
        return loss.detach()

    def _reduce_bucket(self, view: torch.Tensor):
        if view.numel() == 0:
            return
        if self._reduce_fn is not None:
            self._reduce_fn(view)
        elif self.world_size > 1:
            dist = torch.distributed
            self._works.append(dist.all_reduce(view, op=dist.ReduceOp.SUM, group=self.pg, async_op=True))

    def _reduce_writer_bucket(self):
        """Called from the boundary node's backward: every writer gradient kernel has been issued (on the writer streams
        and the side stream); the communication stream picks up exactly these dependencies (event waits only)."""
        from . import model as _model
        cur = torch.cuda.current_stream(self.device)
        comm = self._comm_stream
        comm.wait_stream(cur)
        if _model.PARALLEL_WRITERS:
            for wst in _model._writer_streams(self.device):
                comm.wait_stream(wst)
        for side in ops.side_streams(self.device):
            comm.wait_stream(side)

    def _state_for(self, packed: PackedBatch):
        sig = packed.signature
        st = self._static.pop(sig, None)
        if st is None:
            while len(self._static) >= max(1, self.max_cached_shapes):        # evict the least recently used shape
                old = self._static.pop(next(iter(self._static)))
                if old.get("graph") is not None or old.get("staging") is not None:
                    torch.cuda.synchronize(self.device)                        # nothing of it may still be in flight
                old.clear()
            st = {"batch": DeviceBatch(packed, self.device), "graph": None, "loss": None, "warm": 0}
        self._static[sig] = st                                                 # most recently used last
        return st

    # ---- input pipelining ---------------------------------------------------------------------------------------------
    def prefetch(self, packed: PackedBatch):
        """Starts the host->device copy of a FUTURE step's inputs on the copy stream, into staging buffers, so that it
        overlaps the step that is currently running; the step that consumes it only does a device-to-device copy into
        the static buffers of the captured graph."""
        st = self._state_for(packed)
        stg = st.get("staging")
        if stg is None:
            stg = st["staging"] = {"f": torch.empty_like(st["batch"].fbuf), "i": torch.empty_like(st["batch"].ibuf),
                                   "ready": torch.cuda.Event(), "free": None}
        cs = self._copy_stream
        if stg["free"] is not None:
            cs.wait_event(stg["free"])               # the previous consumer has copied the staging buffers out
        with torch.cuda.stream(cs):
            stg["f"].copy_(packed.fbuf, non_blocking=True)
            stg["i"].copy_(packed.plan.host, non_blocking=True)
            stg["ready"].record(cs)
        st["prefetched"] = packed

    def _load(self, st, packed):
        stg = st.get("staging")
        if stg is not None and st.get("prefetched") is packed:
            cur = torch.cuda.current_stream(self.device)
            cur.wait_event(stg["ready"])
            st["batch"].fbuf.copy_(stg["f"], non_blocking=True)
            st["batch"].ibuf.copy_(stg["i"], non_blocking=True)
            stg["free"] = torch.cuda.Event()
            stg["free"].record(cur)
            st["prefetched"] = None
            st["batch"].refresh_host_fields(packed)
        else:
            st["batch"].load(packed)

    def read_loss_async(self, loss: torch.Tensor):
        """Device->host read of a step's loss without stalling the launch of the next step: returns a handle whose
        `.value()` waits for (only) that copy."""
        buf = self._loss_host[self._loss_slot % len(self._loss_host)]
        self._loss_slot += 1
        buf.copy_(loss.reshape(1), non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        return _LossHandle(buf, ev)

    def compute_gradients(self, packed: PackedBatch, load: bool = True) -> torch.Tensor:
        """H2D (optional) + forward + loss + backward into the flat gradient buffer.  Returns the device loss."""
        st = self._state_for(packed)
        if load:
            self._load(st, packed)
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
        if self.overlap_comm or self._switch is not None:     # exchanged inside compute_gradients / by the update kernel itself
            self._grad_scale = 1.0 / self.world_size
            return
        self._grad_scale = all_reduce_sum_(self.flat.grads, self.world_size, self.pg)

    def optimizer_step(self):
        self.t += 1
        self._lr_now = self._next_lr()
        self._adamw_range(0, self.flat.numel)
        if self.split_update:
            for ev in self._stage_ready:
                ev.record(torch.cuda.current_stream(self.device))

    def _init_planes(self):
        L = _lib.lib()
        self._plane_rows = []                       # (param offset, rows, cols, plane offset)
        total = 0
        for n, p in self.flat.named:
            if p.dim() == 2 and p.shape[0] >= 32 and p.shape[1] >= 32:
                self._plane_rows.append((self.flat.offsets[n], int(p.shape[0]), int(p.shape[1]), total, p))
                total += int(L.grappa_split_bf16_elems(int(p.shape[0]), int(p.shape[1])))
        self._planes = torch.zeros(max(total, 8), dtype=torch.int16, device=self.device)
        self.flat.planes_of = {id(p): ops.WeightPlanes(self._planes, base, r, c) for (_, r, c, base, p) in self._plane_rows}
        self._refresh_planes(0, self.flat.numel)

    def _refresh_planes(self, lo: int, hi: int):
        key = (lo, hi)
        tab = self._plane_tables.get(key)
        if tab is None:
            rows = [(o, r, c, b) for (o, r, c, b, _) in self._plane_rows if lo <= o < hi]
            mt = max([((r + 31) // 32) * ((c + 31) // 32) for (_, r, c, _) in rows], default=0)
            tab = (torch.tensor(rows, dtype=torch.int64, device=self.device).reshape(-1) if rows else None, len(rows), mt)
            self._plane_tables[key] = tab
        if tab[1]:
            _lib.check(_lib.lib().grappa_split_bf16(tab[1], ops._p(tab[0]), tab[2], ops._p(self.flat.params), ops._p(self._planes),
                                                    ops._stream()))

    def restart_warmup(self):
        self._warmup_step = 0 if self.warmup_steps > 0 else None

    def _next_lr(self) -> float:
        """lightning_model.py:136-153: the step counter advances once per optimisation step."""
        if self._warmup_step is None:
            return self.lr
        if self._warmup_step >= self.warmup_steps:
            self._warmup_step = None
            return self.lr
        lr = float(self._warmup_step) / self.warmup_steps * self.lr
        self._warmup_step += 1
        return lr

    def _adamw_range(self, lo: int, hi: int):
        f = self.flat
        if hi <= lo:
            return
        ptr = lambda t: t.data_ptr() + 4 * lo
        if self._switch is not None:       # exchange + update in one kernel; launches that may overlap use different flag slots
            sw = self._switch
            n_ctas = max(1, min(self._dp_ctas, (shard_of(lo, hi, 0, sw.world)[1] // 4 + 2047) // 2048))
            slot = self._slot_of.setdefault((lo, hi), len(self._slot_of))
            pads = (_ctypes.c_uint64 * sw.world)(*sw.pad_ptrs)
            _lib.check(_lib.lib().grappa_dp_exchange_adamw(lo, hi, sw.rank, sw.world, f.grads_mc, f.params_mc, f.params.data_ptr(),
                                                           f.exp_avg.data_ptr(), f.exp_avg_sq.data_ptr(), pads, sw.PAD_WORDS,
                                                           slot * self._dp_ctas * sw.world, n_ctas, self._lr_now, self.betas[0],
                                                           self.betas[1], self.eps, self.wd, self.t, 1.0 / self.world_size,
                                                           self._dp_spin_seconds, ops._stream()))
        else:
            _lib.check(_lib.lib().grappa_adamw(hi - lo, ptr(f.params), ptr(f.grads), ptr(f.exp_avg), ptr(f.exp_avg_sq),
                                               self._lr_now, self.betas[0], self.betas[1], self.eps, self.wd, self.t,
                                               getattr(self, "_grad_scale", 1.0 / self.world_size), ops._stream()))
        ops.WEIGHT_EPOCH[0] += 1         # invalidates lo planes cached for use outside this driver
        if self._planes is not None:     # bf16 planes of the updated weights for the next step's forward / dX GEMMs
            self._refresh_planes(lo, hi)
        elif ops.USE_WEIGHT_LO:          # tf32 lo planes of the updated weights (B operand pre-split, 3xTF32 path)
            _lib.check(_lib.lib().grappa_split_lo(hi - lo, ptr(f.params), ptr(f.params_lo), ops._stream()))

    def exchange_and_update(self):
        """Gradient exchange + AdamW.  With `split_update` the first parameter stage goes on the compute stream and the later
        ones (exchange + AdamW + plane refresh) on the update stream while the next step's forward is computing."""
        if not self.split_update or self.overlap_comm or self.bucket_split in (0, self.flat.numel):
            self.all_reduce_gradients()
            self.optimizer_step()
            return
        dist = torch.distributed
        cur = torch.cuda.current_stream(self.device)
        b = self._stage_bounds
        works = [None] * (len(b) - 1)
        if self.world_size > 1 and self._switch is None:
            for i in range(len(b) - 1):                            # issued in forward order: NCCL completes them in that order
                works[i] = dist.all_reduce(self.flat.grads[b[i]:b[i + 1]], op=dist.ReduceOp.SUM, group=self.pg, async_op=True)
        self._grad_scale = 1.0 / self.world_size
        self.t += 1
        self._lr_now = self._next_lr()
        self._upd_stream.wait_stream(cur)                       # gradients of this step are complete
        if works[0] is not None:
            works[0].wait()
        self._adamw_range(b[0], b[1])
        self._stage_ready[0].record(cur)
        with torch.cuda.stream(self._upd_stream):
            for i in range(1, len(b) - 1):
                if works[i] is not None:
                    works[i].wait()
                self._adamw_range(b[i], b[i + 1])
                self._stage_ready[i].record(self._upd_stream)

    def gather_optimizer_state(self):
        """With the fused NVSwitch exchange every rank updates only its shard of exp_avg / exp_avg_sq; this completes all
        ranks' copies (call before writing a checkpoint or comparing optimiser state across ranks)."""
        if self._switch is None:
            return
        self.synchronize()
        gather_shards_((self.flat.exp_avg, self.flat.exp_avg_sq), list(self._slot_of), self._switch.group)
        torch.cuda.synchronize(self.device)

    def synchronize(self):
        """All parameter updates issued so far have completed on the device."""
        torch.cuda.current_stream(self.device).wait_stream(self._upd_stream)
        torch.cuda.synchronize(self.device)

    def step_from_store(self, store, mol_ids, conf_strategy="max", generator=None) -> torch.Tensor:
        """One optimisation step on a batch collated ON THE DEVICE from a resident `store.MoleculeStore`: the host only
        uploads the segment tables (tens of kB), not the batch."""
        spec = store.prepare(mol_ids, conf_strategy, generator)
        st = self._state_for(spec)
        store.collate_into(spec, st["batch"])
        loss = self.compute_gradients(spec, load=False)
        self.exchange_and_update()
        return loss

    def step(self, packed: PackedBatch, load: bool = True) -> torch.Tensor:
        loss = self.compute_gradients(packed, load=load)
        self.exchange_and_update()
        return loss
