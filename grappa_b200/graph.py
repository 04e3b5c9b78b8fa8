"""
Minimal heterograph container with the accessor surface of a DGL heterograph that the hot path uses
(SURVEY.md §8b "Reads"/"Writes"): `g.ntypes`, `g.num_nodes(t)`, `g.nodes[t].data[...]`, `g.batch_num_nodes(t)`,
`g.edges(etype='n1_edge')`, `g.to(device)`.  `GrappaModel` / `Energy` in this package accept either this container
or a real `dgl.DGLGraph` (they only touch that shared surface).

`batch` / `unbatch` follow the reference's `grappa.utils.dgl_utils.batch` (utils/dgl_utils.py:11-63) and `unbatch`
(:66-87): `idxs` of graph i are offset by the number of atoms of graphs 0..i-1, node data are concatenated per type in
list order, bond edges are offset, per-type `batch_num_nodes` are recorded.  Deliberate difference: the reference
writes the offset `idxs` back into the INPUT graphs (:57-59, `copied_graph = graph`); this implementation leaves its
inputs untouched.

`molecule_graph` follows `Molecule.to_dgl` (data/molecule.py:493-602): idxs int64, bond edges int32 ordered
[all i->j, then all j->i], atomic number one-hot with Z-1 as the hot column.
"""
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

LEVELS = ("n2", "n3", "n4", "n4_improper")
N_SEQ = {"n2": 2, "n3": 3, "n4": 4, "n4_improper": 4}


class _NodeView:
    __slots__ = ("data",)

    def __init__(self, data):
        self.data = data


class _Nodes:
    def __init__(self, g):
        self._g = g

    def __getitem__(self, ntype):
        return _NodeView(self._g._ndata[ntype])


class HeteroGraph:
    def __init__(self, num_nodes: Dict[str, int], edges: Optional[torch.Tensor] = None,
                 batch_num_nodes: Optional[Dict[str, torch.Tensor]] = None):
        """
        num_nodes: {ntype: count}.  edges: (2, E) int32 directed bond edges between 'n1' nodes (src row 0, dst row 1).
        batch_num_nodes: {ntype: int64 tensor (n_graphs,)} (host tensors; they describe segments, not device data).
        """
        self._num_nodes = {k: int(v) for k, v in num_nodes.items()}
        self._ndata: Dict[str, Dict[str, torch.Tensor]] = {nt: {} for nt in self._num_nodes}
        self._edges = edges if edges is not None else torch.zeros((2, 0), dtype=torch.int32)
        if batch_num_nodes is None:
            batch_num_nodes = {nt: torch.tensor([n], dtype=torch.int64) for nt, n in self._num_nodes.items()}
        self._batch_num_nodes = {k: v.to("cpu", torch.int64) for k, v in batch_num_nodes.items()}
        self._plan_cache = {}   # host/device batch plans keyed by kind (filled lazily by grappa_b200.plan)

    # ---- DGL-like surface -------------------------------------------------------------------------------------
    @property
    def ntypes(self) -> List[str]:
        return sorted(self._num_nodes.keys())

    def num_nodes(self, ntype: Optional[str] = None) -> int:
        if ntype is None:
            return sum(self._num_nodes.values())
        return self._num_nodes[ntype]

    def num_edges(self, etype: Optional[str] = None) -> int:
        return int(self._edges.shape[1])

    def edges(self, etype: Optional[str] = None):
        return self._edges[0], self._edges[1]

    @property
    def nodes(self):
        return _Nodes(self)

    def batch_num_nodes(self, ntype: str) -> torch.Tensor:
        return self._batch_num_nodes[ntype]

    @property
    def batch_size(self) -> int:
        return int(len(self._batch_num_nodes["n1"]))

    @property
    def device(self) -> torch.device:
        for d in self._ndata.values():
            for v in d.values():
                return v.device
        return self._edges.device

    def to(self, device, non_blocking: bool = False) -> "HeteroGraph":
        if torch.device(device).type == "cuda" and self.device.type == "cpu" and "plan" not in self._plan_cache \
                and "n1" in self._num_nodes:
            # host -> device (the `g.to(device)` of grappa.py:67): build the batch plan now, from the host-resident index
            # tensors, so that it travels with the graph instead of being derived from device copies (a read-back)
            # (attached to the device copy only: the host graph stays free to be edited)
            try:
                from .plan import build_plan
                fresh_plan = build_plan(self)
            except Exception:
                fresh_plan = None            # malformed graphs raise where the reference raises: in the model's forward
        else:
            fresh_plan = None
        g = HeteroGraph(self._num_nodes, self._edges.to(device, non_blocking=non_blocking), self._batch_num_nodes)
        for nt, d in self._ndata.items():
            for k, v in d.items():
                g._ndata[nt][k] = v.to(device, non_blocking=non_blocking)
        # plans built from host-side index tensors travel with the graph
        g._plan_cache = {k: (p.to(device) if hasattr(p, "to") else p) for k, p in self._plan_cache.items()}
        if fresh_plan is not None:
            g._plan_cache["plan"] = fresh_plan.to(device)
        return g

    def pin_memory(self) -> "HeteroGraph":
        g = HeteroGraph(self._num_nodes, self._edges.pin_memory(), self._batch_num_nodes)
        for nt, d in self._ndata.items():
            for k, v in d.items():
                g._ndata[nt][k] = v.pin_memory()
        g._plan_cache = {k: (p.pin_memory() if hasattr(p, "pin_memory") else p) for k, p in self._plan_cache.items()}
        return g

    def __repr__(self):
        return f"HeteroGraph(num_nodes={self._num_nodes}, num_edges={self.num_edges()}, batch_size={self.batch_size})"


# ---------------------------------------------------------------------------------------------------------------

def molecule_graph(bonds: np.ndarray, angles: np.ndarray, propers: np.ndarray, impropers: np.ndarray,
                   atomic_numbers: Sequence[int], partial_charges: Sequence[float],
                   additional_features: Optional[Dict[str, np.ndarray]] = None,
                   max_element: int = 53) -> HeteroGraph:
    """Builds the single-molecule heterograph with the layout of `Molecule.to_dgl` (data/molecule.py:493-602)."""
    n_atoms = len(atomic_numbers)
    idxs = {
        "n2": torch.as_tensor(np.asarray(bonds, dtype=np.int64).reshape(-1, 2)),
        "n3": torch.as_tensor(np.asarray(angles, dtype=np.int64).reshape(-1, 3)),
        "n4": torch.as_tensor(np.asarray(propers, dtype=np.int64).reshape(-1, 4)),
        "n4_improper": torch.as_tensor(np.asarray(impropers, dtype=np.int64).reshape(-1, 4)),
    }
    b = idxs["n2"].t()
    if b.numel() and int(b.max()) >= n_atoms:
        raise AssertionError(f"Maximal atom index in bonds ({int(b.max())}) must be smaller than the number of atoms ({n_atoms})")
    if len(torch.unique(b.flatten())) != n_atoms:
        raise AssertionError("Every atom must be part of a bond")
    edges = torch.stack((torch.cat((b[0], b[1])), torch.cat((b[1], b[0]))), dim=0).int()
    num_nodes = {"n1": n_atoms, "g": 1}
    num_nodes.update({lvl: int(idxs[lvl].shape[0]) for lvl in LEVELS})
    g = HeteroGraph(num_nodes, edges)
    for lvl in LEVELS:
        g.nodes[lvl].data["idxs"] = idxs[lvl]
    g.nodes["n1"].data["ids"] = torch.arange(n_atoms, dtype=torch.long)
    z = torch.as_tensor(np.asarray(atomic_numbers, dtype=np.int64))
    if int(z.max()) > max_element or int(z.min()) < 1:
        raise ValueError("atomic numbers out of range")
    g.nodes["n1"].data["atomic_number"] = torch.nn.functional.one_hot(z - 1, num_classes=max_element).float()
    g.nodes["n1"].data["partial_charge"] = torch.as_tensor(np.asarray(partial_charges, dtype=np.float32))
    for k, v in (additional_features or {}).items():
        g.nodes["n1"].data[k] = torch.as_tensor(np.asarray(v, dtype=np.float32))
    return g


def batch(graphs: List[HeteroGraph]) -> HeteroGraph:
    """Reference: utils/dgl_utils.py:11-63 (see module docstring for the one deliberate difference)."""
    ntypes = graphs[0].ntypes
    n_atoms = [g.num_nodes("n1") for g in graphs]
    offsets = np.concatenate([[0], np.cumsum(n_atoms[:-1])]).astype(np.int64)
    num_confs = None
    if "xyz" in graphs[0].nodes["n1"].data:
        num_confs = graphs[0].nodes["n1"].data["xyz"].shape[1]
        for g in graphs:
            c = g.nodes["n1"].data["xyz"].shape[1]
            if c != num_confs:
                raise ValueError(f"All graphs must have the same number of conformations but found {num_confs} and {c}")
    num_nodes = {nt: sum(g.num_nodes(nt) for g in graphs) for nt in ntypes}
    bnn = {nt: torch.tensor([g.num_nodes(nt) for g in graphs], dtype=torch.int64) for nt in ntypes}
    edges = torch.cat([g._edges + int(o) for g, o in zip(graphs, offsets)], dim=1)
    out = HeteroGraph(num_nodes, edges, bnn)
    for nt in ntypes:
        for key in graphs[0].nodes[nt].data.keys():
            parts = [g.nodes[nt].data[key] for g in graphs]
            if key == "idxs" and nt in LEVELS:
                parts = [p + int(o) for p, o in zip(parts, offsets)]
            out.nodes[nt].data[key] = torch.cat(parts, dim=0)
    return out


def unbatch(g: HeteroGraph) -> List[HeteroGraph]:
    """Reference: utils/dgl_utils.py:66-87 incl. removal of dummy conformations (:90-125)."""
    n = g.batch_size
    if n == 1:
        return [g]
    offs = {nt: 0 for nt in g.ntypes}
    src, dst = g.edges()
    outs = []
    for i in range(n):
        cnt = {nt: int(g.batch_num_nodes(nt)[i]) for nt in g.ntypes}
        lo = offs["n1"]
        m = (src >= lo) & (src < lo + cnt["n1"])
        sg = HeteroGraph(cnt, torch.stack((src[m] - lo, dst[m] - lo), dim=0))
        for nt in g.ntypes:
            for key, v in g.nodes[nt].data.items():
                part = v[offs[nt]:offs[nt] + cnt[nt]]
                if key == "idxs" and nt in LEVELS:
                    part = part - lo
                sg.nodes[nt].data[key] = part
        for nt in g.ntypes:
            offs[nt] += cnt[nt]
        outs.append(delete_dummy_confs(sg))
    return outs


def delete_dummy_confs(g: HeteroGraph) -> HeteroGraph:
    """Reference: utils/dgl_utils.py:90-125."""
    if "is_dummy" not in g.nodes["g"].data:
        return g
    keep = g.nodes["g"].data["is_dummy"][0] == 0
    if bool(keep.all()):
        return g
    gd = g.nodes["g"].data
    gd["is_dummy"] = gd["is_dummy"][:, keep]
    n1 = g.nodes["n1"].data
    n1["xyz"] = n1["xyz"][:, keep, :]
    for k in list(gd.keys()):
        if "energy" in k:
            gd[k] = gd[k][:, keep]
    for k in list(n1.keys()):
        if "gradient" in k:
            n1[k] = n1[k][:, keep, :]
    for lvl in LEVELS:
        if lvl in g.ntypes:
            d = g.nodes[lvl].data
            for k in list(d.keys()):
                if "energy" in k:
                    d[k] = d[k][:, keep]
    return g


def conf_indices(confs_present: int, num_confs: int, generator: Optional[torch.Generator] = None) -> torch.Tensor:
    """Which stored conformations fill the `num_confs` slots of a batch (utils/dgl_utils.py:147-162): all of them in
    order when the counts agree, a random subset (`randperm[:num_confs]`) when there are more, all followed by repeats
    of the last one (flagged as dummies) when there are fewer."""
    if confs_present == num_confs:
        return torch.arange(num_confs)
    if confs_present > num_confs:
        return torch.randperm(confs_present, generator=generator)[:num_confs]
    return torch.cat((torch.arange(confs_present), torch.full((num_confs - confs_present,), confs_present - 1, dtype=torch.long)))


def set_number_confs(g: HeteroGraph, num_confs: int, generator: Optional[torch.Generator] = None,
                     conf_idxs: Optional[torch.Tensor] = None) -> HeteroGraph:
    """Reference: utils/dgl_utils.py:137-176.  Returns a NEW graph (the input is not mutated) whose conformation axis has
    length `num_confs`; `g.nodes['g'].data['is_dummy']` marks padded slots.  `conf_idxs` overrides the random choice."""
    if "xyz" not in g.nodes["n1"].data:
        return g
    present = g.nodes["n1"].data["xyz"].shape[1]
    if conf_idxs is None:
        conf_idxs = conf_indices(present, num_confs, generator)
    out = HeteroGraph(dict(g._num_nodes), g._edges, {k: v.clone() for k, v in g._batch_num_nodes.items()})
    for nt in g.ntypes:
        out.nodes[nt].data.update(g.nodes[nt].data)
    gd, n1 = out.nodes["g"].data, out.nodes["n1"].data
    dummy = torch.zeros((1, num_confs), dtype=torch.float32)
    if present < num_confs:
        dummy[0, present:] = 1.0
    gd["is_dummy"] = dummy
    n1["xyz"] = n1["xyz"][:, conf_idxs]
    for k in list(gd.keys()):
        if "energy" in k:
            gd[k] = gd[k][:, conf_idxs]
            if torch.isnan(gd[k]).any():
                raise RuntimeError(f"Found nan in {k} after setting number of conformations to {num_confs}")
    for k in list(n1.keys()):
        if "gradient" in k:
            n1[k] = n1[k][:, conf_idxs]
    return out


def get_bond_edges(g):
    """(src, dst) of the directed bond edges for either container type."""
    try:
        return g.edges(etype="n1_edge")
    except TypeError:
        return g.edges()
