"""
Device-resident molecule store + device-side collate.

The reference collates on the host, per step, inside DataLoader workers: it picks the batch's conformation count
(`conf_strategy`), sub-samples or pads every molecule's conformations (`set_number_confs`, is_dummy flags), offsets
the `idxs`, concatenates everything (`dgl_utils.batch`), pickles the batched graph to the trainer and copies it to
the GPU tensor by tensor (data/graph_data_loader.py:23-73, utils/dgl_utils.py:11-63,137-176).

Here every molecule is packed ONCE (its fp32 fields and its own int32 batch plan) into two buffers that stay in HBM
(180 GB per GPU holds any of the reference's datasets many times over).  A step's batch is then assembled on the GPU
by two gather launches (`grappa_collate_float`, `grappa_collate_int`): all per-molecule plan arrays compose under
batching by adding an offset, so the batch plan is a concatenation with per-segment addends.  Per step the host only
does O(n_mol) numpy bookkeeping (offsets, conformation choice with the reference's rule) and uploads a segment table
of a few tens of kB instead of the whole batch.

The result is bit-identical to `PackedBatch(batch([set_number_confs(m, C, conf_idxs=...) for m in mols]))` loaded
into a `DeviceBatch` (tests/test_gpu_collate.py), so everything downstream (CUDA-graph replay included) is unchanged.
"""
from typing import Dict, List, Optional, Sequence, Tuple, Union

import numpy as np
import torch

from . import _lib
from .batching import DeviceBatch, _FLOAT_FIELDS
from .graph import LEVELS, N_SEQ, HeteroGraph, conf_indices, delete_dummy_confs
from .plan import ENERGY_SIZE_KEYS, BatchPlan, build_plan

_SIZE_KEYS = ENERGY_SIZE_KEYS
_SUM_KEYS = ("n_blocks", "n_words")


def _pad4(n: int) -> int:
    return (int(n) + 3) // 4 * 4


class CollatedSpec:
    """Shapes of one collated batch (what `PackedBatch` exposes to `DeviceBatch` / `TrainStep`) plus the upload tables."""

    def __init__(self):
        self.num_nodes: Dict[str, int] = {}
        self.batch_num_nodes: Dict[str, torch.Tensor] = {}
        self.n_edges = 0
        self.layout: Dict[Tuple[str, str], Tuple[int, Tuple[int, ...]]] = {}
        self.plan: Optional[BatchPlan] = None
        self.n_float = 0
        self.n_int = 0
        self.edges = None
        self.seg64: Optional[torch.Tensor] = None     # host (n_int_seg + n_float_seg, 8) int64
        self.tab32: Optional[torch.Tensor] = None     # host int32: [conformation table | literals]
        self.n_int_seg = 0
        self.n_float_seg = 0
        self.lit_off = 0
        self.conf_idxs: Optional[torch.Tensor] = None  # (n_mol, C) the conformations chosen per molecule

    @property
    def signature(self):
        return (tuple(sorted(self.num_nodes.items())), tuple(sorted((k, v) for k, v in self.layout.items())),
                tuple(sorted(self.plan.offsets.items())), tuple(sorted(self.plan.energy_sizes.items())))

    @property
    def h2d_bytes(self) -> int:
        return int(self.seg64.numel()) * 8 + int(self.tab32.numel()) * 4


class MoleculeStore:
    def __init__(self, molecules: Sequence[HeteroGraph], device, feature_names: Optional[List[str]] = None,
                 energy_parts: Optional[int] = None):
        """`energy_parts`: warps per CTA of the energy kernel for batches collated from this store (every molecule's
        plan is compiled for it).  Default: 16 (training batches have few CTAs, hence many short parts: plan.build_plan),
        4 when a molecule spans several atom blocks (un-compacted accumulators: parts x block atoms rows of shared memory)."""
        if energy_parts is None:
            from .plan import ENERGY_BLOCK_ATOMS, ENERGY_PARTS
            energy_parts = ENERGY_PARTS if max(int(g.num_nodes("n1")) for g in molecules) > ENERGY_BLOCK_ATOMS else 16
        self.energy_parts = int(energy_parts)
        if len(molecules) == 0:
            raise ValueError("empty molecule store")
        self.device = torch.device(device)
        self.n = len(molecules)
        mols = [delete_dummy_confs(_shallow(g)) for g in molecules]
        g0 = mols[0]
        n1 = g0.nodes["n1"].data
        self.feature_names = feature_names if feature_names is not None else [
            k for k in n1.keys() if k not in ("xyz", "gradient_ref", "ids", "h", "gradient") and torch.is_floating_point(n1[k])]
        self.ntypes = list(g0.ntypes)
        self.float_fields: List[Tuple[str, str]] = [("n1", k) for k in self.feature_names] + [
            (nt, k) for nt, k in _FLOAT_FIELDS if k != "is_dummy" and k in g0.nodes[nt].data]
        self.inner: Dict[Tuple[str, str], Tuple[int, ...]] = {}       # trailing shape after (rows[, C])
        self.has_confs: Dict[Tuple[str, str], bool] = {}
        for nt, k in self.float_fields:
            t = g0.nodes[nt].data[k]
            confs = k in ("xyz", "gradient_ref", "energy_ref")
            self.has_confs[(nt, k)] = confs
            self.inner[(nt, k)] = tuple(t.shape[2:]) if confs else tuple(t.shape[1:])

        # ---- pack ------------------------------------------------------------------------------------------------
        M = self.n
        self.N = np.zeros(M, np.int64)
        self.E = np.zeros(M, np.int64)
        self.C = np.zeros(M, np.int64)
        self.T = {lvl: np.zeros(M, np.int64) for lvl in LEVELS}
        self.sizes = {k: np.zeros(M, np.int64) for k in _SIZE_KEYS}
        self.int_names: Optional[List[str]] = None
        self.i_off: Dict[str, np.ndarray] = {}
        self.i_len: Dict[str, np.ndarray] = {}
        self.f_off: Dict[Tuple[str, str], np.ndarray] = {f: np.zeros(M, np.int64) for f in self.float_fields}
        self.levels_present = None
        fparts, iparts = [], []
        fbase = ibase = 0
        for m, g in enumerate(mols):
            if list(g.ntypes) != self.ntypes:
                raise ValueError("all molecules of a store must have the same node types")
            plan = build_plan(g, energy=True, n_parts=energy_parts)
            if plan.n_mol != 1:
                raise ValueError("store entries must be single molecules")
            if self.int_names is None:
                self.int_names = list(plan.offsets.keys())
                self.levels_present = plan.levels_present
                for name in self.int_names:
                    self.i_off[name] = np.zeros(M, np.int64)
                    self.i_len[name] = np.zeros(M, np.int64)
            self.N[m], self.E[m] = plan.n_atoms, plan.n_edges
            for lvl in LEVELS:
                self.T[lvl][m] = plan.T[lvl]
            for k in _SIZE_KEYS:
                self.sizes[k][m] = plan.energy_sizes[k]
            for name in self.int_names:
                s, n = plan.offsets[name]
                self.i_off[name][m] = ibase + s
                self.i_len[name][m] = n
            iparts.append(plan.host)
            ibase += int(plan.host.numel())
            self.C[m] = g.nodes["n1"].data["xyz"].shape[1] if "xyz" in g.nodes["n1"].data else 0
            for f in self.float_fields:
                t = g.nodes[f[0]].data[f[1]].reshape(-1).float()
                self.f_off[f][m] = fbase
                fparts.append(t)
                pad = _pad4(t.numel()) - t.numel()
                if pad:
                    fparts.append(torch.zeros(pad))
                fbase += _pad4(t.numel())
        self.fstore = torch.cat(fparts).to(self.device) if fparts else torch.zeros(4, device=self.device)
        self.istore = torch.cat(iparts).to(self.device)
        self._ring = []          # pinned staging buffers: (seg64, tab32, event)
        self._ring_pos = 0
        self._dev_seg = None
        self._dev_tab = None

    @property
    def nbytes(self) -> int:
        return int(self.fstore.numel()) * 4 + int(self.istore.numel()) * 4

    # ---- host bookkeeping for one batch ----------------------------------------------------------------------------
    def n_confs_for(self, mol_ids: np.ndarray, conf_strategy: Union[str, int]) -> int:
        """data/graph_data_loader.py:52-66."""
        c = self.C[mol_ids]
        if isinstance(conf_strategy, (int, np.integer)):
            return int(min(int(conf_strategy), c.max()))
        if conf_strategy == "min":
            return int(c.min())
        if conf_strategy in ("max", "all"):
            return int(c.max())
        if conf_strategy == "mean":
            return int(np.mean(c))
        raise ValueError(f"Unknown conf_strategy: {conf_strategy}")

    def prepare(self, mol_ids: Sequence[int], conf_strategy: Union[str, int] = "max",
                generator: Optional[torch.Generator] = None, conf_idxs: Optional[torch.Tensor] = None) -> CollatedSpec:
        ids = np.asarray(mol_ids, dtype=np.int64).reshape(-1)
        if len(ids) == 0:
            raise ValueError("empty batch")
        if ids.min() < 0 or ids.max() >= self.n:
            raise IndexError(f"molecule id out of range [0,{self.n})")
        B = len(ids)
        N, E = self.N[ids], self.E[ids]
        T = {lvl: self.T[lvl][ids] for lvl in LEVELS}
        sz = {k: self.sizes[k][ids] for k in _SIZE_KEYS}
        cum = lambda a: np.concatenate([[0], np.cumsum(a)]).astype(np.int64)
        Ac, Ec = cum(N), cum(E)
        Tc = {lvl: cum(T[lvl]) for lvl in LEVELS}
        Bc, Wc = cum(sz["n_blocks"]), cum(sz["n_words"])
        nA, nE = int(Ac[-1]), int(Ec[-1])
        has_xyz = ("n1", "xyz") in self.f_off
        C = self.n_confs_for(ids, conf_strategy) if has_xyz else 0
        if conf_idxs is not None:
            C = int(conf_idxs.shape[1])

        spec = CollatedSpec()
        spec.num_nodes = {nt: (nA if nt == "n1" else B if nt == "g" else int(Tc[nt][-1])) for nt in self.ntypes}
        spec.batch_num_nodes = {nt: torch.from_numpy(N.copy() if nt == "n1" else np.ones(B, np.int64) if nt == "g"
                                                      else T[nt].copy()) for nt in self.ntypes}
        spec.n_edges = nE

        # ---- int layout (same section order and padding as plan.build_plan) ------------------------------------------
        plan = BatchPlan()
        plan.n_atoms, plan.n_mol, plan.n_edges = nA, B, nE
        plan.T = {lvl: int(Tc[lvl][-1]) for lvl in LEVELS}
        plan.levels_present = self.levels_present
        plan.mol_atom_counts = N.copy()
        plan.energy_sizes = {k: int(sz[k].sum()) if k in _SUM_KEYS else int(sz[k].max()) for k in _SIZE_KEYS}
        if int(sz["n_parts"].min()) != plan.energy_sizes["n_parts"]:
            raise ValueError("all molecules of a store must be planned with the same number of energy parts")
        T_all = sum(plan.T.values())
        lengths = {}
        for lvl in LEVELS:
            lengths["idx_" + lvl] = N_SEQ[lvl] * plan.T[lvl]
            lengths["inc_ptr_" + lvl] = nA + 1
            lengths["inc_tok_" + lvl] = N_SEQ[lvl] * plan.T[lvl]
            lengths["mol_off_" + lvl] = B + 1
            lengths["tab_idx_" + lvl] = N_SEQ[lvl] * plan.T[lvl]
            lengths["tab_ptr_" + lvl] = 2 * nA + 1
            lengths["tab_tok_" + lvl] = N_SEQ[lvl] * plan.T[lvl]
        lengths.update(in_ptr=nA + 1, out_ptr=nA + 1, in_nbr=nE, out_nbr=nE, e_blk=12 * int(Bc[-1]),
                       e_prog=int(Wc[-1]), e_mol_toff=4 * (B + 1), tuple_mol=T_all, mol_atom_off=B + 1)
        total = 0
        for name in self.int_names:
            plan.offsets[name] = (total, int(lengths[name]))
            total += _pad4(lengths[name])
        spec.n_int = max(total, 4)
        spec.plan = plan

        # ---- int segments ----------------------------------------------------------------------------------------
        zeros = np.zeros(B, np.int64)
        mol = np.arange(B, dtype=np.int64)
        rows = []
        lits: List[np.ndarray] = []
        lit_pos = 0

        def seg(name, dst_rel, length, mode, a0=zeros, a1=zeros, a2=zeros, a3=zeros, src=None):
            base = plan.offsets[name][0]
            s = self.i_off[name][ids] if src is None else src
            rows.append(np.stack([s, base + dst_rel, length, np.full(B, mode, np.int64), a0, a1, a2, a3], axis=1))

        def literal(name, values):
            nonlocal lit_pos
            v = np.asarray(values, dtype=np.int32)
            rows.append(np.array([[lit_pos, plan.offsets[name][0], len(v), 5, 0, 0, 0, 0]], dtype=np.int64))
            lits.append(v)
            lit_pos += len(v)

        tuple_base = 0
        for name in self.int_names:
            ln = self.i_len[name][ids]
            if name.startswith("idx_"):
                lvl = name[4:]
                seg(name, N_SEQ[lvl] * Tc[lvl][:-1], ln, 0, Ac[:-1])
            elif name.startswith("inc_ptr_"):
                lvl = name[8:]
                seg(name, Ac[:-1], ln, 0, N_SEQ[lvl] * Tc[lvl][:-1])
            elif name.startswith("inc_tok_"):
                lvl = name[8:]
                seg(name, N_SEQ[lvl] * Tc[lvl][:-1], ln, 0, N_SEQ[lvl] * Tc[lvl][:-1])
            elif name.startswith("tab_idx_"):
                lvl = name[8:]
                seg(name, N_SEQ[lvl] * Tc[lvl][:-1], ln, 0, 2 * Ac[:-1])
            elif name.startswith("tab_ptr_"):
                lvl = name[8:]
                seg(name, 2 * Ac[:-1], ln, 0, N_SEQ[lvl] * Tc[lvl][:-1])
            elif name.startswith("tab_tok_"):
                lvl = name[8:]
                seg(name, N_SEQ[lvl] * Tc[lvl][:-1], ln, 0, N_SEQ[lvl] * Tc[lvl][:-1])
            elif name in ("in_ptr", "out_ptr"):
                seg(name, Ac[:-1], ln, 0, Ec[:-1])
            elif name in ("in_nbr", "out_nbr"):
                seg(name, Ec[:-1], ln, 0, Ac[:-1])
            elif name == "e_blk":
                seg(name, 12 * Bc[:-1], ln, 1, Ac[:-1], mol, Wc[:-1])
            elif name == "e_prog":
                seg(name, Wc[:-1], ln, 0)          # every program word is relative to its block / molecule
            elif name == "e_mol_toff":
                literal(name, np.concatenate([Tc[lvl] for lvl in LEVELS]))
            elif name == "tuple_mol":
                for lvl in LEVELS:
                    seg(name, tuple_base + Tc[lvl][:-1], T[lvl], 4, mol, src=zeros)
                    tuple_base += plan.T[lvl]
            elif name == "mol_atom_off":
                literal(name, Ac)
            elif name.startswith("mol_off_"):
                literal(name, Tc[name[8:]])
            else:
                raise AssertionError(f"no collate rule for plan array {name}")
        int_seg = np.concatenate(rows, axis=0)
        int_seg = int_seg[int_seg[:, 2] > 0]

        # ---- float layout + segments -----------------------------------------------------------------------------
        if conf_idxs is None and has_xyz:
            conf_idxs = torch.stack([conf_indices(int(c), C, generator) for c in self.C[ids]])
        spec.conf_idxs = conf_idxs
        frows = []
        ftotal = 0
        fields = list(self.float_fields) + ([("g", "is_dummy")] if has_xyz else [])
        for f in fields:
            nt, k = f
            if k == "is_dummy":
                shape = (B, C)
                frows.append(np.stack([zeros, ftotal + mol * C, zeros, zeros, self.C[ids], np.full(B, C, np.int64), zeros - 1,
                                       np.ones(B, np.int64)], axis=1))
            else:
                r = N if nt == "n1" else np.ones(B, np.int64)
                rc = cum(r)
                inner = int(np.prod(self.inner[f])) if len(self.inner[f]) else 1
                if self.has_confs[f]:
                    shape = (int(rc[-1]), C) + self.inner[f]
                    frows.append(np.stack([self.f_off[f][ids], ftotal + rc[:-1] * C * inner, r, np.full(B, inner, np.int64),
                                           self.C[ids], np.full(B, C, np.int64), mol * C, zeros], axis=1))
                else:
                    shape = (int(rc[-1]),) + self.inner[f]
                    frows.append(np.stack([self.f_off[f][ids], ftotal + rc[:-1] * inner, r, np.full(B, inner, np.int64),
                                           np.ones(B, np.int64), np.ones(B, np.int64), zeros - 1, zeros], axis=1))
            spec.layout[f] = (ftotal, tuple(int(x) for x in shape))
            ftotal += _pad4(int(np.prod(shape)))
        spec.n_float = max(ftotal, 4)
        float_seg = np.concatenate(frows, axis=0) if frows else np.zeros((0, 8), np.int64)

        spec.n_int_seg, spec.n_float_seg = len(int_seg), len(float_seg)
        spec.seg64 = torch.from_numpy(np.ascontiguousarray(np.concatenate([int_seg, float_seg], axis=0)))
        conf_flat = conf_idxs.to(torch.int32).reshape(-1).numpy() if conf_idxs is not None else np.zeros(0, np.int32)
        spec.lit_off = _pad4(len(conf_flat))
        tab = np.zeros(spec.lit_off + _pad4(lit_pos) + 4, np.int32)
        tab[:len(conf_flat)] = conf_flat
        if lits:
            tab[spec.lit_off:spec.lit_off + lit_pos] = np.concatenate(lits)
        spec.tab32 = torch.from_numpy(tab)
        return spec

    # ---- device side -------------------------------------------------------------------------------------------
    def _stage(self, spec: CollatedSpec):
        """Copies the tables into pinned staging (ring of 4, guarded by events) and on to the device."""
        n64, n32 = int(spec.seg64.numel()), int(spec.tab32.numel())
        if len(self._ring) < 4:
            self._ring.append([None, None, None])
        slot = self._ring[self._ring_pos % len(self._ring)]
        self._ring_pos += 1
        if slot[2] is not None:
            slot[2].synchronize()
        if slot[0] is None or slot[0].numel() < n64:
            slot[0] = torch.empty(max(n64, 1024) * 2, dtype=torch.int64).pin_memory()
        if slot[1] is None or slot[1].numel() < n32:
            slot[1] = torch.empty(max(n32, 1024) * 2, dtype=torch.int32).pin_memory()
        slot[0][:n64].copy_(spec.seg64.reshape(-1))
        slot[1][:n32].copy_(spec.tab32)
        if self._dev_seg is None or self._dev_seg.numel() < n64:
            self._dev_seg = torch.empty(max(n64, 1024) * 2, dtype=torch.int64, device=self.device)
        if self._dev_tab is None or self._dev_tab.numel() < n32:
            self._dev_tab = torch.empty(max(n32, 1024) * 2, dtype=torch.int32, device=self.device)
        self._dev_seg[:n64].copy_(slot[0][:n64], non_blocking=True)
        self._dev_tab[:n32].copy_(slot[1][:n32], non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        slot[2] = ev

    def collate_into(self, spec: CollatedSpec, batch: DeviceBatch) -> HeteroGraph:
        """Uploads the segment tables and assembles the batch in `batch`'s static buffers (current stream)."""
        if batch.signature != spec.signature:
            raise ValueError("batch shapes differ from the static device buffers")
        self._stage(spec)
        batch.refresh_host_fields(spec)
        L = _lib.lib()
        stream = torch.cuda.current_stream(self.device).cuda_stream
        seg, tab = self._dev_seg.data_ptr(), self._dev_tab.data_ptr()
        _lib.check(L.grappa_collate_int(spec.n_int_seg, seg, self.istore.data_ptr(), tab + 4 * spec.lit_off,
                                        batch.ibuf.data_ptr(), stream))
        _lib.check(L.grappa_collate_float(spec.n_float_seg, seg + 64 * spec.n_int_seg, tab, self.fstore.data_ptr(),
                                          batch.fbuf.data_ptr(), stream))
        return batch.graph

    def collate(self, mol_ids: Sequence[int], conf_strategy: Union[str, int] = "max",
                generator: Optional[torch.Generator] = None, conf_idxs: Optional[torch.Tensor] = None) -> DeviceBatch:
        """One-off convenience: fresh device buffers for this batch."""
        spec = self.prepare(mol_ids, conf_strategy, generator, conf_idxs)
        batch = DeviceBatch(spec, self.device)
        self.collate_into(spec, batch)
        return batch


def _shallow(g: HeteroGraph) -> HeteroGraph:
    out = HeteroGraph(dict(g._num_nodes), g._edges, {k: v.clone() for k, v in g._batch_num_nodes.items()})
    for nt in g.ntypes:
        out.nodes[nt].data.update(g.nodes[nt].data)
    return out
