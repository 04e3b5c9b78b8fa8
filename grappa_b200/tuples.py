"""
Interaction-tuple enumeration: bonds -> angles, proper torsions, improper triples.

Vectorised (numpy) restatement of the layout produced by the reference's
`grappa.utils.tuple_indices.get_idx_tuples` (utils/tuple_indices.py:8-70) and `get_torsions`
(utils/tuple_indices.py:160-246).  The ORDER of the rows is part of the contract ("defines
tuple layout that must be bit-exact"), so the sort keys below reproduce the reference's loop nest:

  for atom1 in first-appearance order of the bond list            (dict insertion order, :35, :83-91)
    for atom2 in sorted(neigh[atom1])                             (:37)
      for atom3 in sorted(neigh[atom2]), atom3 != atom1           (:38-40)
        angle (atom1, atom2, atom3)            if atom1 < atom3   (:42-44)
        for atom4 in sorted(neigh[atom3]) while atom4 < atom1, atom4 != atom2:
          proper (atom4, atom3, atom2, atom1)                     (:48-56)

This is host-side index work (int64), not a GPU kernel.  `enumerate_tuples` runs the loop nest in C++ through the
C ABI (`grappa_enumerate_tuples`, csrc/tuples.cu); `enumerate_tuples_numpy` is an independent sort-based derivation
of the same layout kept as a cross-check for the tests.
"""
import ctypes
from typing import Dict, Sequence, Tuple

import numpy as np


def _adjacency(bonds: np.ndarray, n_atoms: int) -> Tuple[np.ndarray, np.ndarray]:
    """CSR adjacency (indptr, indices) with every neighbour list sorted ascending."""
    src = np.concatenate([bonds[:, 0], bonds[:, 1]])
    dst = np.concatenate([bonds[:, 1], bonds[:, 0]])
    order = np.lexsort((dst, src))
    src, dst = src[order], dst[order]
    indptr = np.zeros(n_atoms + 1, dtype=np.int64)
    np.add.at(indptr, src + 1, 1)
    indptr = np.cumsum(indptr)
    return indptr, dst.astype(np.int64)


def _first_appearance_rank(bonds: np.ndarray, n_atoms: int) -> np.ndarray:
    """rank[a] = position of atom a in the order in which atoms first appear in the flattened bond list."""
    flat = bonds.reshape(-1)
    _, first = np.unique(flat, return_index=True)
    atoms = np.unique(flat)
    rank = np.full(n_atoms, np.iinfo(np.int64).max, dtype=np.int64)
    rank[atoms[np.argsort(first, kind="stable")]] = np.arange(len(atoms))
    return rank


def _expand(indptr: np.ndarray, indices: np.ndarray, at: np.ndarray):
    """For every entry i of `at`, list the neighbours of at[i]. Returns (row index into `at`, neighbour)."""
    deg = indptr[at + 1] - indptr[at]
    row = np.repeat(np.arange(len(at)), deg)
    start = np.repeat(indptr[at], deg)
    within = np.arange(len(row)) - np.repeat(np.cumsum(deg) - deg, deg)
    return row, indices[start + within]


def enumerate_tuples(bonds: Sequence[Sequence[int]], n_atoms: int = None) -> Dict[str, np.ndarray]:
    """
    bonds: (n_bonds, 2) atom indices in 0..n_atoms-1.
    Returns {'bonds': (n_bonds,2) with bond[0] < bond[1], 'angles': (n_angles,3), 'propers': (n_propers,4)} int64,
    rows in the reference's order (see module docstring).  Raises ValueError on a self-bond.
    """
    from . import _lib
    bonds = np.ascontiguousarray(np.asarray(bonds, dtype=np.int64).reshape(-1, 2))
    if n_atoms is None:
        n_atoms = int(bonds.max()) + 1 if len(bonds) else 0
    L = _lib.lib()
    ptr = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    counts = np.zeros(2, np.int64)
    rc = L.grappa_enumerate_tuples(len(bonds), ptr(bonds), int(n_atoms), ptr(counts), None, None)
    if rc != 0:
        raise ValueError(L.grappa_last_error().decode())
    angles = np.empty((int(counts[0]), 3), np.int64)
    propers = np.empty((int(counts[1]), 4), np.int64)
    rc = L.grappa_enumerate_tuples(len(bonds), ptr(bonds), int(n_atoms), ptr(counts), ptr(angles), ptr(propers))
    if rc != 0:
        raise ValueError(L.grappa_last_error().decode())
    return {"bonds": np.sort(bonds, axis=1), "angles": angles, "propers": propers}


def enumerate_tuples_numpy(bonds: Sequence[Sequence[int]], n_atoms: int = None) -> Dict[str, np.ndarray]:
    """Sort-based derivation of the same rows (no loop nest): cross-check for `enumerate_tuples`."""
    bonds = np.asarray(bonds, dtype=np.int64).reshape(-1, 2)
    if n_atoms is None:
        n_atoms = int(bonds.max()) + 1 if len(bonds) else 0
    if len(bonds) == 0:
        return {"bonds": bonds, "angles": np.zeros((0, 3), np.int64), "propers": np.zeros((0, 4), np.int64)}
    if np.any(bonds[:, 0] == bonds[:, 1]):
        raise ValueError("self-bond encountered")

    indptr, indices = _adjacency(bonds, n_atoms)
    rank = _first_appearance_rank(bonds, n_atoms)

    # directed edges a1 -> a2 (every atom with its sorted neighbours)
    a1 = np.repeat(np.arange(n_atoms, dtype=np.int64), indptr[1:] - indptr[:-1])
    a2 = indices
    # paths a1 - a2 - a3
    r, a3 = _expand(indptr, indices, a2)
    p1, p2 = a1[r], a2[r]
    keep = p1 != a3
    p1, p2, a3 = p1[keep], p2[keep], a3[keep]

    ang = p1 < a3
    angles = np.stack([p1[ang], p2[ang], a3[ang]], axis=1)
    order = np.lexsort((angles[:, 2], angles[:, 1], rank[angles[:, 0]]))
    angles = angles[order]

    # paths a1 - a2 - a3 - a4 with a4 < a1 and a4 != a2
    r, a4 = _expand(indptr, indices, a3)
    q1, q2, q3 = p1[r], p2[r], a3[r]
    keep = (a4 < q1) & (a4 != q2)
    q1, q2, q3, a4 = q1[keep], q2[keep], q3[keep], a4[keep]
    order = np.lexsort((a4, q3, q2, rank[q1]))
    propers = np.stack([a4[order], q3[order], q2[order], q1[order]], axis=1)

    return {"bonds": np.sort(bonds, axis=1), "angles": angles.astype(np.int64), "propers": propers.astype(np.int64)}


def improper_rows(centres: Sequence[int], neighbours: Sequence[Sequence[int]], central_pos: int = 2) -> np.ndarray:
    """
    Three cyclic versions per improper centre with the central atom at `central_pos`
    (reference: utils/tuple_indices.py:213-238, constants.IMPROPER_CENTRAL_IDX = 2).
    neighbours[i] are the three outer atoms of centres[i] in the order they appear in the caller's torsion.
    Returns (3*n, 4) int64.
    """
    centres = np.asarray(centres, dtype=np.int64).reshape(-1)
    nb = np.asarray(neighbours, dtype=np.int64).reshape(-1, 3)
    if len(centres) == 0:
        return np.zeros((0, 4), np.int64)
    out = np.empty((len(centres), 3, 4), dtype=np.int64)
    outer_pos = [p for p in range(4) if p != central_pos]
    for v, cyc in enumerate(((0, 1, 2), (1, 2, 0), (2, 0, 1))):
        out[:, v, central_pos] = centres
        for j, p in enumerate(outer_pos):
            out[:, v, p] = nb[:, cyc[j]]
    return out.reshape(-1, 4)
