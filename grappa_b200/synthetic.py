"""
Synthetic molecular graphs for the BASELINE.json configurations.  There is no network for datasets,
so tests and bench.py build inputs of the right SHAPE here: capped alanine dipeptide (22 atoms), an alternating
ALA/PHE polymer without caps (3 residues = 40 atoms, 1000 residues = 15,000 atoms), random-walk geometries with bond
lengths in 1.0-1.5 Angstrom and N(0, 0.1^2) displacement per conformation, features drawn as described in §8d.

The heterograph layout is the one `Molecule.to_dgl` produces (data/molecule.py:493-602); tuple enumeration is
`grappa_b200.tuples` (bit-exact vs utils/tuple_indices.py, checked in tests/test_tuples.py).
"""
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from .graph import HeteroGraph, batch, molecule_graph
from .tuples import enumerate_tuples, improper_rows

# element by name prefix
_Z = {"H": 1, "C": 6, "N": 7, "O": 8}

_ALA = dict(
    names=["N", "H", "CA", "HA", "CB", "HB1", "HB2", "HB3", "C", "O"],
    bonds=[("N", "H"), ("N", "CA"), ("CA", "HA"), ("CA", "CB"), ("CB", "HB1"), ("CB", "HB2"), ("CB", "HB3"),
           ("CA", "C"), ("C", "O")],
    improper_centres=[],
)
_PHE = dict(
    names=["N", "H", "CA", "HA", "CB", "HB2", "HB3", "CG", "CD1", "HD1", "CE1", "HE1", "CZ", "HZ", "CE2", "HE2",
           "CD2", "HD2", "C", "O"],
    bonds=[("N", "H"), ("N", "CA"), ("CA", "HA"), ("CA", "CB"), ("CB", "HB2"), ("CB", "HB3"), ("CB", "CG"),
           ("CG", "CD1"), ("CD1", "HD1"), ("CD1", "CE1"), ("CE1", "HE1"), ("CE1", "CZ"), ("CZ", "HZ"), ("CZ", "CE2"),
           ("CE2", "HE2"), ("CE2", "CD2"), ("CD2", "HD2"), ("CD2", "CG"), ("CA", "C"), ("C", "O")],
    improper_centres=["CG", "CD1", "CE1", "CZ", "CE2", "CD2"],
)
_RES = {"A": _ALA, "F": _PHE}


def template_ace_ala_nme() -> Dict[str, np.ndarray]:
    """Capped alanine dipeptide, 22 atoms, 21 bonds, impropers at the 4 sp2 centres (BASELINE cfg 1)."""
    names = ["HH31", "CH3", "HH32", "HH33", "C", "O", "N", "H", "CA", "HA", "CB", "HB1", "HB2", "HB3", "C", "O",
             "N", "H", "CH3", "HH31", "HH32", "HH33"]
    bonds = [(0, 1), (1, 2), (1, 3), (1, 4), (4, 5), (4, 6), (6, 7), (6, 8), (8, 9), (8, 10), (10, 11), (10, 12),
             (10, 13), (8, 14), (14, 15), (14, 16), (16, 17), (16, 18), (18, 19), (18, 20), (18, 21)]
    z = [_Z[n[0]] for n in names]
    return dict(bonds=np.array(bonds, np.int64), atomic_numbers=np.array(z, np.int64),
                improper_centres=np.array([4, 6, 14, 16], np.int64))


def template_polymer(sequence: str) -> Dict[str, np.ndarray]:
    """Uncapped peptide chain; `sequence` over {'A' (ALA, 10 atoms), 'F' (PHE, 20 atoms)}.
    Improper centres: aromatic ring carbons and every backbone carbonyl C that has three neighbours."""
    bonds: List = []
    z: List[int] = []
    centres: List[int] = []
    prev_c = None
    off = 0
    for r in sequence:
        res = _RES[r]
        idx = {n: off + i for i, n in enumerate(res["names"])}
        z += [_Z[n[0]] for n in res["names"]]
        bonds += [(idx[a], idx[b]) for a, b in res["bonds"]]
        centres += [idx[c] for c in res["improper_centres"]]
        if prev_c is not None:
            bonds.append((prev_c, idx["N"]))
            centres.append(prev_c)
        prev_c = idx["C"]
        off += len(res["names"])
    return dict(bonds=np.array(bonds, np.int64), atomic_numbers=np.array(z, np.int64),
                improper_centres=np.array(sorted(centres), np.int64))


def alternating_sequence(n_res: int, start: str = "A") -> str:
    other = "F" if start == "A" else "A"
    return "".join(start if i % 2 == 0 else other for i in range(n_res))


def _neighbour_lists(bonds: np.ndarray, n: int) -> List[List[int]]:
    nb: List[List[int]] = [[] for _ in range(n)]
    for a, b in bonds.tolist():
        nb[a].append(b)
        nb[b].append(a)
    return [sorted(x) for x in nb]


def random_walk_geometry(bonds: np.ndarray, n_atoms: int, rng: np.random.Generator) -> np.ndarray:
    """Places atoms by BFS over the bond graph: child = parent + L * u, L ~ U(1.0, 1.5), with u chosen so that all
    angles at the parent lie in roughly (70, 160) degrees (no collinear or folded-back triples)."""
    nb = _neighbour_lists(bonds, n_atoms)
    pos = np.zeros((n_atoms, 3))
    placed = np.zeros(n_atoms, bool)
    placed[0] = True
    queue = [0]
    while queue:
        p = queue.pop(0)
        dirs = [(pos[q] - pos[p]) / np.linalg.norm(pos[q] - pos[p]) for q in nb[p] if placed[q]]
        for c in nb[p]:
            if placed[c]:
                continue
            for _ in range(200):
                u = rng.normal(size=3)
                u /= np.linalg.norm(u)
                if all(-0.94 < float(u @ d) < 0.34 for d in dirs):
                    break
            pos[c] = pos[p] + rng.uniform(1.0, 1.5) * u
            placed[c] = True
            dirs.append(u)
            queue.append(c)
    assert placed.all(), "bond graph is not connected"
    return pos


def make_molecule(template: Dict[str, np.ndarray], seed: int = 0, n_confs: Optional[int] = 50,
                  with_reference_data: bool = False, random_ring_encoding: bool = False) -> HeteroGraph:
    bonds = template["bonds"]
    z = template["atomic_numbers"]
    n = len(z)
    rng = np.random.default_rng(seed)
    tup = enumerate_tuples(bonds, n)
    nb = _neighbour_lists(bonds, n)
    centres = template["improper_centres"]
    imp = improper_rows(centres, [nb[c] for c in centres.tolist()])
    deg = np.array([len(x) for x in nb])
    degree = np.zeros((n, 6), np.float32)
    degree[np.arange(n), np.clip(deg, 1, 6) - 1] = 1.0
    ring = np.zeros((n, 7), np.float32)
    if random_ring_encoding:
        ring[np.arange(n), rng.integers(0, 7, size=n)] = (rng.random(n) < 0.3).astype(np.float32)
    q = rng.normal(0.0, 0.3, size=n).astype(np.float32)
    g = molecule_graph(tup["bonds"], tup["angles"], tup["propers"], imp, z, q,
                       additional_features={"ring_encoding": ring, "degree": degree})
    if n_confs is not None:
        base = random_walk_geometry(bonds, n, rng)
        xyz = base[:, None, :] + rng.normal(0.0, 0.1, size=(n, n_confs, 3))
        g.nodes["n1"].data["xyz"] = torch.as_tensor(xyz.astype(np.float32))
        g.nodes["g"].data["is_dummy"] = torch.zeros((1, n_confs), dtype=torch.float32)
        if with_reference_data:
            g.nodes["g"].data["energy_ref"] = torch.as_tensor(rng.normal(size=(1, n_confs)).astype(np.float32))
            g.nodes["n1"].data["gradient_ref"] = torch.as_tensor(rng.normal(size=(n, n_confs, 3)).astype(np.float32))
    return g


# ---- the BASELINE.json configurations ----------------------------------------------------------------------------

_TRIMERS = ("AFA", "FAA", "AAF")  # all 40 atoms / 40 bonds; the graph differs with the position of the ring


def cfg1_graph(n_confs: int = 50, seed: int = 0) -> HeteroGraph:
    """22-atom capped alanine dipeptide."""
    return make_molecule(template_ace_ala_nme(), seed=seed, n_confs=n_confs)


def trimer_molecule(seed: int, n_confs: int = 50, with_reference_data: bool = True) -> HeteroGraph:
    return make_molecule(template_polymer(_TRIMERS[seed % 3]), seed=seed, n_confs=n_confs,
                         with_reference_data=with_reference_data)


def cfg2_graphs(n_mol: int = 32, n_confs: int = 50, first_seed: int = 0) -> List[HeteroGraph]:
    """SPICE-like training batch: n_mol 40-atom molecules x n_confs conformations (rank r uses first_seed = 32 r)."""
    return [trimer_molecule(first_seed + i, n_confs) for i in range(n_mol)]


def cfg2_batch(n_mol: int = 32, n_confs: int = 50, first_seed: int = 0) -> HeteroGraph:
    return batch(cfg2_graphs(n_mol, n_confs, first_seed))


def cfg3_graph(n_res: int = 1000, seed: int = 0, n_confs: Optional[int] = None) -> HeteroGraph:
    """1000-residue alternating polymer: 15,000 atoms, parameters only (what Grappa.predict runs)."""
    return make_molecule(template_polymer(alternating_sequence(n_res)), seed=seed, n_confs=n_confs)


def cfg4_batch(n_mol: int = 1000, n_confs: int = 500, seed: int = 0) -> HeteroGraph:
    """Energy sweep: n_mol copies of the 40-atom molecule topology with independent geometries."""
    tmpl = template_polymer("AFA")
    first = make_molecule(tmpl, seed=seed, n_confs=None)
    n = first.num_nodes("n1")
    rng = np.random.default_rng(seed + 1)
    base = random_walk_geometry(tmpl["bonds"], n, rng)
    graphs = []
    for i in range(n_mol):
        g = HeteroGraph(first._num_nodes, first._edges)
        for nt in first.ntypes:
            g._ndata[nt] = dict(first._ndata[nt])
        graphs.append(g)
    out = batch(graphs)
    xyz = np.tile(base, (n_mol, 1))[:, None, :].astype(np.float32) + \
        rng.normal(0.0, 0.1, size=(n * n_mol, n_confs, 3)).astype(np.float32)
    out.nodes["n1"].data["xyz"] = torch.as_tensor(xyz)
    return out
