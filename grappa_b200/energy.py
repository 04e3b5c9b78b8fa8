"""
`Energy`: drop-in for `grappa.models.energy.Energy` (models/energy.py:74-195) on the fused energy/force kernel.

Same constructor keywords, same graph reads (`xyz`, `idxs`, `k`/`eq` + suffix) and writes (`energy`, `energy_<term>`,
`gradient`, optional `gradient_<term>`, per-tuple `x` / `energy`), same exceptions.  Differences, all deliberate:
  - forces are analytic (no autograd.grad, no second-order graph): `gradient` is differentiable w.r.t. the MM
    parameters through the closed-form VJP kernel, which is what the training loss needs; it is not differentiable
    w.r.t. xyz (the reference's double backward w.r.t. coordinates is never used: xyz is a leaf input).
  - the dihedral noise of internal_coordinates.py:178-180 is not added (deterministic kernel, explicit guards).
  - xyz.requires_grad is left untouched (the reference flips it in place, energy.py:127).
  - `materialize_tuple_terms=False` skips writing the (T, C) per-tuple coordinates/energies.
"""
import torch

from . import ops
from .plan import ensure_energy, get_plan

_LEVELS = ("n2", "n3", "n4", "n4_improper")


class Energy(torch.nn.Module):
    def __init__(self, terms: list = ["bond", "angle", "torsion", "improper"], suffix: str = "", offset_torsion: bool = False,
                 write_suffix=None, gradients: bool = True, gradient_contributions: bool = False,
                 materialize_tuple_terms: bool = True):
        super().__init__()
        if not isinstance(terms, list):
            raise ValueError("terms must be a list")
        if offset_torsion:
            raise NotImplementedError("grappa_b200: offset_torsion=True is not part of the accelerated path")
        self.offset_torsion = offset_torsion
        self.suffix = suffix
        self.write_suffix = write_suffix if write_suffix is not None else suffix
        self.gradients = gradients
        self.gradient_contributions = gradient_contributions
        self.materialize_tuple_terms = materialize_tuple_terms
        self.TERM_TO_LEVEL = {"bond": "n2", "angle": "n3", "proper": "n4", "torsion": "n4", "improper": "n4_improper"}
        self.LEVEL_TO_TERM = {v: k for k, v in self.TERM_TO_LEVEL.items()}
        self.LEVEL_TO_TERM["n4"] = "proper"
        self.terms = [self.TERM_TO_LEVEL[t] for t in terms]

    def forward(self, g):
        if self.gradient_contributions and not self.gradients:
            raise ValueError("Gradient contributions cannot be calculated if gradients are not enabled.")
        n1 = g.nodes["n1"].data
        if "xyz" not in n1.keys():
            raise ValueError("xyz coordinates must be stored in g.nodes['n1'].data['xyz']")
        xyz = n1["xyz"]
        plan = ensure_energy(g, get_plan(g))
        params = {}
        mask = 0
        for term in self.terms:
            if term not in g.ntypes:
                raise ValueError(f"term {term} not in g.ntypes")
            d = g.nodes[term].data
            if "k" + self.suffix not in d.keys():
                raise RuntimeError(f"{term} has no k{self.suffix} attribute")
            params[term] = (d["k" + self.suffix], d.get("eq" + self.suffix) if term in ("n2", "n3") else None)
            mask |= 1 << _LEVELS.index(term)

        def kq(level):
            return params.get(level, (None, None))

        k2, eq2 = kq("n2")
        k3, eq3 = kq("n3")
        k4, k4i = kq("n4")[0], kq("n4_improper")[0]
        outs = ops.EnergyFn.apply(k2, eq2, k3, eq3, k4, k4i, xyz, plan, mask, self.materialize_tuple_terms)
        energy, grad, term_energy = outs[0], outs[1], outs[2]
        ws = self.write_suffix
        for li, level in enumerate(_LEVELS):
            if level not in self.terms:
                continue
            g.nodes["g"].data["energy_" + ws + self.LEVEL_TO_TERM[level]] = term_energy[li]
            if self.materialize_tuple_terms:
                g.nodes[level].data["x"] = outs[3 + li]
                g.nodes[level].data["energy" + ws] = outs[7 + li]
        g.nodes["g"].data["energy" + ws] = energy
        if self.gradients:
            g.nodes["n1"].data["gradient" + ws] = grad
            if self.gradient_contributions:
                for li, level in enumerate(_LEVELS):
                    if level not in self.terms:
                        continue
                    with torch.no_grad():
                        o = ops.EnergyFn.apply(k2, eq2, k3, eq3, k4, k4i, xyz, plan, 1 << li, False)
                    g.nodes["n1"].data["gradient_" + ws + self.LEVEL_TO_TERM[level]] = o[1]
        return g
