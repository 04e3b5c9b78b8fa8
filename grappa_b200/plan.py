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

    # -- access -------------------------------------------------------------------------------------------------
    def view(self, name: str) -> torch.Tensor:
        s, n = self.offsets[name]
        return self.dev[s:s + n]

    def ptr(self, name: str) -> int:
        s, n = self.offsets[name]
        return self.dev.data_ptr() + 4 * s

    def host_view(self, name: str) -> np.ndarray:
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


def build_plan(g, block_atoms: int = ENERGY_BLOCK_ATOMS, n_parts: Optional[int] = None) -> BatchPlan:
    L = _lib.lib()
    p = BatchPlan()
    n_atoms = int(g.num_nodes("n1"))
    p.n_atoms = n_atoms
    bnn = g.batch_num_nodes("n1").to("cpu", torch.int64).numpy()
    p.n_mol = int(len(bnn))
    p.mol_atom_counts = bnn
    mol_atom_off = np.zeros(p.n_mol + 1, np.int64)
    np.cumsum(bnn, out=mol_atom_off[1:])
    if int(mol_atom_off[-1]) != n_atoms:
        raise AssertionError("batch_num_nodes('n1') does not sum to the number of atoms")

    arrays: Dict[str, np.ndarray] = {}
    idx64: Dict[str, np.ndarray] = {}
    present = []
    for lvl in LEVELS:
        if lvl in g.ntypes:
            present.append(lvl)
            ix = g.nodes[lvl].data["idxs"]
            ix = ix.detach().to("cpu", torch.int64).contiguous().numpy().reshape(-1, N_SEQ[lvl])
            if ix.size:
                mx, mn = int(ix.max()), int(ix.min())
                # reference: models/grappa.py:179-185 (AssertionError on out-of-range idxs)
                assert mx < n_atoms and mn >= 0, (
                    f'Encountered idxs up to {mx} at the level g.nodes[{lvl}].data["idxs"], but there are only '
                    f'{n_atoms} atom-level-nodes in the graph')
        else:
            ix = np.zeros((0, N_SEQ[lvl]), np.int64)
        idx64[lvl] = ix
        p.T[lvl] = int(ix.shape[0])
        arrays["idx_" + lvl] = ix.astype(np.int32).reshape(-1)
        # atom -> token rows (row = t * n_seq + i), stable in row order
        inc_ptr = np.zeros(n_atoms + 1, np.int32)
        inc_tok = np.zeros(ix.size, np.int32)
        _lib.check(L.grappa_csr_build(ix.size, n_atoms, _np_ptr(np.ascontiguousarray(ix.reshape(-1))), _np_ptr(inc_ptr),
                                      _np_ptr(inc_tok)))
        arrays["inc_ptr_" + lvl] = inc_ptr
        arrays["inc_tok_" + lvl] = inc_tok
        if lvl in PE_PATTERN:
            # table row of a token = 2 * atom + pe(position): composes under batching with the offset 2 * atom_offset
            keys = np.ascontiguousarray((2 * ix + np.asarray(PE_PATTERN[lvl], np.int64)[None, :]).reshape(-1))
            tab_ptr = np.zeros(2 * n_atoms + 1, np.int32)
            tab_tok = np.zeros(ix.size, np.int32)
            _lib.check(L.grappa_csr_build(ix.size, 2 * n_atoms, _np_ptr(keys), _np_ptr(tab_ptr), _np_ptr(tab_tok)))
            arrays["tab_idx_" + lvl] = keys.astype(np.int32)
            arrays["tab_ptr_" + lvl] = tab_ptr
            arrays["tab_tok_" + lvl] = tab_tok
    p.levels_present = tuple(present)

    # bond edges -> CSR by destination / by source
    src, dst = get_bond_edges(g)
    src = src.detach().to("cpu", torch.int64).contiguous().numpy()
    dst = dst.detach().to("cpu", torch.int64).contiguous().numpy()
    p.n_edges = int(len(src))
    for name, key, other in (("in", dst, src), ("out", src, dst)):
        ptr = np.zeros(n_atoms + 1, np.int32)
        order = np.zeros(len(key), np.int32)
        _lib.check(L.grappa_csr_build(len(key), n_atoms, _np_ptr(key), _np_ptr(ptr), _np_ptr(order)))
        arrays[name + "_ptr"] = ptr
        arrays[name + "_nbr"] = other[order].astype(np.int32)

    # energy plan
    sizes = np.zeros(12, np.int64)
    args_idx = []
    keep = []
    for lvl in LEVELS:
        a = np.ascontiguousarray(idx64[lvl].reshape(-1))
        args_idx += [_np_ptr(a), p.T[lvl]]
        keep.append(a)   # keep alive until the second call
    if n_parts is None:
        n_parts = int(os.environ.get("GRAPPA_ENERGY_PARTS", ENERGY_PARTS))
    _lib.check(L.grappa_energy_plan_build(n_atoms, p.n_mol, _np_ptr(mol_atom_off), *args_idx, block_atoms, n_parts,
                                          _np_ptr(sizes), None, None, None, None))
    nB, nW = int(sizes[0]), int(sizes[1])
    blk = np.zeros(nB * 12, np.int32)
    prog = np.zeros(nW, np.int32)
    mol_toff = np.zeros(4 * (p.n_mol + 1), np.int32)
    tuple_mol = np.zeros(sum(p.T.values()), np.int32)
    _lib.check(L.grappa_energy_plan_build(n_atoms, p.n_mol, _np_ptr(mol_atom_off), *args_idx, block_atoms, n_parts,
                                          _np_ptr(sizes), _np_ptr(blk), _np_ptr(prog), _np_ptr(mol_toff), _np_ptr(tuple_mol)))
    del keep
    p.energy_sizes = dict(zip(ENERGY_SIZE_KEYS, (int(v) for v in sizes)))
    arrays.update(e_blk=blk, e_prog=prog, e_mol_toff=mol_toff, tuple_mol=tuple_mol)

    # per-molecule offsets (int32) for the fused loss / segmenting
    arrays["mol_atom_off"] = mol_atom_off.astype(np.int32)
    for lvl in LEVELS:
        off = np.zeros(p.n_mol + 1, np.int64)
        if lvl in g.ntypes:
            np.cumsum(g.batch_num_nodes(lvl).to("cpu", torch.int64).numpy(), out=off[1:])
        arrays["mol_off_" + lvl] = off.astype(np.int32)

    # pack (16-byte aligned sections)
    total = 0
    for k, a in arrays.items():
        p.offsets[k] = (total, int(a.size))
        total += (int(a.size) + 3) // 4 * 4
    buf = np.zeros(max(total, 4), np.int32)
    for k, a in arrays.items():
        s, n = p.offsets[k]
        buf[s:s + n] = a
    p.host = torch.from_numpy(buf)
    return p


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
