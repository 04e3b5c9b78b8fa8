"""
Host batch plan: everything index-shaped that the kernels need, built once per batch on the
host from the graph's int64 `idxs`, the bond edge list and the per-type `batch_num_nodes`, packed into ONE int32
buffer and moved to the device with a single copy:

  - idxs narrowed to int32 per level                         (gather / VJP kernels)
  - CSR of the directed bond edges by destination and by source   (graph attention fwd / bwd)
  - atom -> token-row incidence CSR per level                (pull-style backward of the tuple gather)
  - (atom, pe) table row per token + its CSR for the levels with a positional-encoding column
                                                             (token de-duplication of the first LayerNorm + QKV)
  - the energy kernel's atom-block table and per-block tuple programs     (grappa_energy_plan_build)
  - per-molecule offsets of atoms / proper / improper rows   (segmenting, fused loss)

The reference has no such object: it re-derives the same information on every call through advanced indexing and
DGL's segment bookkeeping (models/interaction_parameters.py:184, models/internal_coordinates.py:31-94,
models/energy.py:69).  Index bounds are validated here (reference: models/grappa.py:179-185 does it with a device sync).
"""
import ctypes
import os
from typing import Dict, Optional

import numpy as np
import torch

from . import _lib
from .graph import LEVELS, N_SEQ, get_bond_edges

ENERGY_BLOCK_ATOMS = 64
ENERGY_PARTS = 4          # warps per CTA of the energy kernel = parts of a block's tuple program
# grappa_energy_plan_build's sizes_out[12]; under batching the first two add up, the others are maxima
ENERGY_SIZE_KEYS = ("n_blocks", "n_words", "max_words", "max_slots", "max_atoms", "any_multi", "n_parts",
                    "max_t2", "max_t3", "max_t4", "max_t4i", "any_nonlocal")
# positional-encoding column of the default architecture per level (perm_equiv_transformer.py:189-199); bonds have none.
# The (atom, pe) token tables below are built for these patterns; a model with another encoding does not use them.
PE_PATTERN = {"n3": (0, 1, 0), "n4": (0, 1, 1, 0), "n4_improper": (0, 1, 1, 0)}


def _np_ptr(a: np.ndarray):
    return a.ctypes.data_as(ctypes.c_void_p)


class BatchPlan:
    """Immutable after construction.  `host` is a packed int32 numpy buffer, `dev` its device copy (or None)."""

    def __init__(self):
        self.n_atoms = 0
        self.n_mol = 0
        self.T: Dict[str, int] = {}
        self.levels_present = ()
        self.n_edges = 0
        self.energy_sizes = None
        self.offsets: Dict[str, tuple] = {}    # name -> (start, length) in int32 elements
        self.host: Optional[torch.Tensor] = None
        self.dev: Optional[torch.Tensor] = None
        self.mol_atom_counts = None
        self.ext: Optional["BatchPlan"] = None  # energy sections added after the fact (ensure_energy)

    # -- access -------------------------------------------------------------------------------------------------
    def view(self, name: str) -> torch.Tensor:
        if name not in self.offsets and self.ext is not None:
            return self.ext.view(name)
        s, n = self.offsets[name]
        return self.dev[s:s + n]

    def ptr(self, name: str) -> int:
        if name not in self.offsets and self.ext is not None:
            return self.ext.ptr(name)
        s, n = self.offsets[name]
        return self.dev.data_ptr() + 4 * s

    def host_view(self, name: str) -> np.ndarray:
        if name not in self.offsets and self.ext is not None:
            return self.ext.host_view(name)
        s, n = self.offsets[name]
        return self.host.numpy()[s:s + n]

    def to(self, device, non_blocking: bool = False) -> "BatchPlan":
        device = torch.device(device)
        if device.type != "cuda":
            return self
        if self.dev is not None and self.dev.device == device:
            return self
        out = BatchPlan()
        out.__dict__.update(self.__dict__)
        out.dev = self.host.to(device, non_blocking=non_blocking)
        if self.ext is not None:
            out.ext = self.ext.to(device, non_blocking=non_blocking)
        return out

    def pin_memory(self) -> "BatchPlan":
        out = BatchPlan()
        out.__dict__.update(self.__dict__)
        out.host = self.host.pin_memory()
        return out

    def energy_sizes_array(self) -> np.ndarray:
        """The plan's sizes in the order grappa_energy_fwd expects (host int64[12])."""
        a = getattr(self, "_es_arr", None)
        if a is None or self._es_src is not self.energy_sizes:
            a = np.asarray([self.energy_sizes[k] for k in ENERGY_SIZE_KEYS], np.int64)
            self._es_arr, self._es_src = a, self.energy_sizes
        return a

    @property
    def h2d_bytes(self) -> int:
        return int(self.host.numel()) * 4


# section order of grappa_plan_build (csrc/plan.cu)
def _section_names():
    names = []
    for lvl in LEVELS:
        names += [f"idx_{lvl}", f"inc_ptr_{lvl}", f"inc_tok_{lvl}", f"tab_idx_{lvl}", f"tab_ptr_{lvl}", f"tab_tok_{lvl}"]
    names += ["in_ptr", "in_nbr", "out_ptr", "out_nbr", "e_blk", "e_prog", "e_mol_toff", "tuple_mol", "mol_atom_off"]
    names += [f"mol_off_{lvl}" for lvl in LEVELS]
    return names


SECTION_NAMES = _section_names()
_ENERGY_SECTIONS = ("e_blk", "e_prog", "e_mol_toff", "tuple_mol")


def _host_i64(t) -> np.ndarray:
    return np.ascontiguousarray(t.detach().to("cpu", torch.int64).contiguous().numpy())


def build_plan(g, block_atoms: int = ENERGY_BLOCK_ATOMS, n_parts: Optional[int] = None, energy: Optional[bool] = None,
               pin: bool = False) -> BatchPlan:
    """One call into grappa_plan_build.  `energy`: include the energy kernel's plan (default: iff the graph carries
    conformations, i.e. `xyz`; Energy.forward adds it lazily otherwise, so parameter-only inference does not pay for it).
    `pin`: the packed buffer is written straight into pinned host memory."""
    L = _lib.lib()
    p = BatchPlan()
    n_atoms = int(g.num_nodes("n1"))
    p.n_atoms = n_atoms
    bnn = _host_i64(g.batch_num_nodes("n1"))
    p.n_mol = int(len(bnn))
    p.mol_atom_counts = bnn
    mol_atom_off = np.zeros(p.n_mol + 1, np.int64)
    np.cumsum(bnn, out=mol_atom_off[1:])
    if int(mol_atom_off[-1]) != n_atoms:
        raise AssertionError("batch_num_nodes('n1') does not sum to the number of atoms")
    if energy is None:
        energy = "xyz" in g.nodes["n1"].data
    if n_parts is None:
        n_parts = int(os.environ.get("GRAPPA_ENERGY_PARTS", 0))
    if n_parts <= 0:
        # warps per CTA of the energy kernel: 4 is best once the grid fills the machine (cfg 4: 16,000 CTAs); a training
        # batch has few (molecule, conformation-chunk) CTAs, each latency-bound, so it is cut into more, shorter parts
        n_parts = ENERGY_PARTS
        if energy and "xyz" in g.nodes["n1"].data:
            ctas = p.n_mol * ((int(g.nodes["n1"].data["xyz"].shape[1]) + 31) // 32)
            n_parts = ENERGY_PARTS if ctas >= 600 else (8 if ctas >= 150 else 16)
            if len(bnn) and int(bnn.max()) > block_atoms:
                n_parts = ENERGY_PARTS      # molecules spanning blocks use un-compacted accumulators (parts x block atoms rows)

    keep = []
    idx_ptrs = (ctypes.c_void_p * 4)()
    cnt_ptrs = (ctypes.c_void_p * 4)()
    pe_ptrs = (ctypes.c_void_p * 4)()
    T = (ctypes.c_int64 * 4)()
    present = []
    for i, lvl in enumerate(LEVELS):
        if lvl in g.ntypes:
            present.append(lvl)
            ix = _host_i64(g.nodes[lvl].data["idxs"]).reshape(-1, N_SEQ[lvl])
            cnt = _host_i64(g.batch_num_nodes(lvl))
            keep += [ix, cnt]
            idx_ptrs[i] = ix.ctypes.data if ix.size else None
            cnt_ptrs[i] = cnt.ctypes.data
            T[i] = ix.shape[0]
        else:
            idx_ptrs[i], cnt_ptrs[i], T[i] = None, None, 0
        p.T[lvl] = int(T[i])
        if lvl in PE_PATTERN:
            pat = np.asarray(PE_PATTERN[lvl], np.int32)
            keep.append(pat)
            pe_ptrs[i] = pat.ctypes.data
        else:
            pe_ptrs[i] = None
    p.levels_present = tuple(present)
    src, dst = get_bond_edges(g)
    src, dst = _host_i64(src), _host_i64(dst)
    p.n_edges = int(len(src))

    n_out = ctypes.c_int64(0)
    sections = np.zeros(2 * len(SECTION_NAMES), np.int64)
    sizes = np.zeros(12, np.int64)
    handle = ctypes.c_void_p(None)
    rc = L.grappa_plan_build(n_atoms, p.n_mol, _np_ptr(mol_atom_off), idx_ptrs, T, cnt_ptrs, pe_ptrs, p.n_edges,
                             _np_ptr(src) if p.n_edges else None, _np_ptr(dst) if p.n_edges else None,
                             1 if energy else 0, block_atoms, n_parts, ctypes.byref(n_out), _np_ptr(sections), _np_ptr(sizes),
                             ctypes.byref(handle))
    if rc == 4:       # reference: models/grappa.py:179-185 (AssertionError on out-of-range idxs)
        msg = L.grappa_last_error().decode()
        for i, lvl in enumerate(LEVELS):
            msg = msg.replace(f"at the level {i},", f'at the level g.nodes[{lvl}].data["idxs"],')
        raise AssertionError(msg)
    _lib.check(rc)
    try:
        host = torch.empty(int(n_out.value), dtype=torch.int32, pin_memory=bool(pin and torch.cuda.is_available()))
        _lib.check(L.grappa_plan_fill(handle, host.data_ptr(), int(n_out.value)))
    finally:
        L.grappa_plan_free(handle)
    for i, name in enumerate(SECTION_NAMES):
        if name in _ENERGY_SECTIONS and not energy:
            continue
        if name.startswith("tab_") and name.split("_", 2)[2] not in PE_PATTERN:
            continue
        p.offsets[name] = (int(sections[2 * i]), int(sections[2 * i + 1]))
    p.energy_sizes = dict(zip(ENERGY_SIZE_KEYS, (int(v) for v in sizes))) if energy else None
    p.host = host
    del keep
    return p


def build_energy_ext(g, plan: BatchPlan, block_atoms: int = ENERGY_BLOCK_ATOMS, n_parts: Optional[int] = None) -> BatchPlan:
    """The energy kernel's plan alone, for a graph whose batch plan was built without it (parameter-only inference that
    later evaluates energies): its own small packed buffer, attached to `plan` as `plan.ext`."""
    L = _lib.lib()
    if n_parts is None:
        n_parts = int(os.environ.get("GRAPPA_ENERGY_PARTS", ENERGY_PARTS))
    mol_atom_off = np.zeros(plan.n_mol + 1, np.int64)
    np.cumsum(np.asarray(plan.mol_atom_counts, np.int64), out=mol_atom_off[1:])
    args, keep = [], []
    for lvl in LEVELS:
        if lvl in g.ntypes:
            a = _host_i64(g.nodes[lvl].data["idxs"]).reshape(-1)
        else:
            a = np.zeros(0, np.int64)
        keep.append(a)
        args += [_np_ptr(a), plan.T[lvl]]
    sizes = np.zeros(12, np.int64)
    _lib.check(L.grappa_energy_plan_build(plan.n_atoms, plan.n_mol, _np_ptr(mol_atom_off), *args, block_atoms, n_parts,
                                          _np_ptr(sizes), None, None, None, None))
    arrays = dict(e_blk=np.zeros(int(sizes[0]) * 12, np.int32), e_prog=np.zeros(int(sizes[1]), np.int32),
                  e_mol_toff=np.zeros(4 * (plan.n_mol + 1), np.int32), tuple_mol=np.zeros(sum(plan.T.values()), np.int32))
    _lib.check(L.grappa_energy_plan_build(plan.n_atoms, plan.n_mol, _np_ptr(mol_atom_off), *args, block_atoms, n_parts,
                                          _np_ptr(sizes), _np_ptr(arrays["e_blk"]), _np_ptr(arrays["e_prog"]),
                                          _np_ptr(arrays["e_mol_toff"]), _np_ptr(arrays["tuple_mol"])))
    ext = BatchPlan()
    total = 0
    for k, a in arrays.items():
        ext.offsets[k] = (total, int(a.size))
        total += (int(a.size) + 3) // 4 * 4
    buf = np.zeros(max(total, 4), np.int32)
    for k, a in arrays.items():
        s0, n = ext.offsets[k]
        buf[s0:s0 + n] = a
    ext.host = torch.from_numpy(buf)
    ext.energy_sizes = dict(zip(ENERGY_SIZE_KEYS, (int(v) for v in sizes)))
    return ext


def ensure_energy(g, plan: BatchPlan) -> BatchPlan:
    """Makes sure `plan` can serve the energy kernel (Energy.forward calls this)."""
    if plan.energy_sizes is not None:
        return plan
    ext = build_energy_ext(g, plan)
    if plan.dev is not None:
        ext = ext.to(plan.dev.device)
    plan.ext = ext
    plan.energy_sizes = ext.energy_sizes
    return plan


def get_plan(g, device=None) -> BatchPlan:
    """Plan cached on the graph object; moved to `device` (the graph's device by default) on first use."""
    cache = getattr(g, "_plan_cache", None)
    if cache is None:
        cache = {}
        try:
            g._plan_cache = cache
        except AttributeError:
            pass
    plan = cache.get("plan")
    if plan is None:
        plan = build_plan(g)
    if device is None:
        device = g.nodes["n1"].data[next(iter(g.nodes["n1"].data))].device
    plan = plan.to(device)
    cache["plan"] = plan
    return plan
