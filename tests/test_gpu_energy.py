"""K4 / K5 parity: fused energy + analytic forces and the parameter VJP against the CPU oracle (which is pinned to
the reference's autograd path by tests/test_oracle_golden.py), plus size-independent properties at full size."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import grappa_oracle as O
from grappa_b200 import ops, synthetic as S
from grappa_b200.energy import Energy
from grappa_b200.graph import batch
from grappa_b200.plan import build_plan, get_plan

DEV = "cuda"
LEVELS = ("n2", "n3", "n4", "n4_improper")


def random_params(g, seed=0, P4=3, P4i=2):
    gen = torch.Generator().manual_seed(seed)
    T = {l: g.num_nodes(l) for l in LEVELS}
    return {
        "n2": {"k": 500 + 300 * torch.rand(T["n2"], generator=gen), "eq": 1.0 + 0.5 * torch.rand(T["n2"], generator=gen)},
        "n3": {"k": 60 + 80 * torch.rand(T["n3"], generator=gen), "eq": 1.5 + 0.8 * torch.rand(T["n3"], generator=gen)},
        "n4": {"k": torch.randn(T["n4"], P4, generator=gen)},
        "n4_improper": {"k": torch.randn(T["n4_improper"], P4i, generator=gen) * 3},
    }


def write_params(g, params, device):
    for l in LEVELS:
        for key, v in params[l].items():
            g.nodes[l].data[key] = v.to(device)


def oracle_energy(g, params, create_graph=False):
    n1, src, dst, idxs, counts = O.graph_inputs(g)
    return O.energy_forward(n1["xyz"], idxs, params, counts, create_graph=create_graph)


def close(a, b, rtol=1e-4, atol=0.0, msg=""):
    a = a.detach().cpu().double().numpy()
    b = b.detach().cpu().double().numpy()
    np.testing.assert_allclose(a, b, rtol=rtol, atol=atol, err_msg=msg)


@pytest.mark.parametrize("case", ["cfg1", "batch3", "multiblock", "one_conf"])
def test_energy_forward_matches_oracle(case, monkeypatch):
    if case == "cfg1":
        g = S.cfg1_graph(n_confs=50, seed=0)
    elif case == "batch3":
        g = batch([S.trimer_molecule(s, 7) for s in (0, 1, 2)])
    elif case == "multiblock":
        g = batch([S.make_molecule(S.template_polymer("AFAFAFAFAF"), seed=2, n_confs=5), S.trimer_molecule(4, 5)])
    else:
        g = S.make_molecule(S.template_polymer(S.alternating_sequence(40)), seed=3, n_confs=1)
    params = random_params(g)
    ref = oracle_energy(g, params)
    gd = g.to(DEV)
    write_params(gd, params, DEV)
    if case == "multiblock":
        assert get_plan(gd).energy_sizes["any_multi"] == 1
    gd = Energy(terms=["bond", "angle", "proper", "improper"])(gd)
    escale = float(ref["energy"].abs().max())
    close(gd.nodes["g"].data["energy"], ref["energy"], atol=1e-5 * escale, msg="total energy")
    for l, term in zip(LEVELS, ("bond", "angle", "proper", "improper")):
        close(gd.nodes["g"].data["energy_" + term], ref["term_energy"][l], atol=1e-5 * escale, msg="term energy " + term)
        close(gd.nodes[l].data["x"], ref["x"][l], atol=2e-6, msg="internal coordinate " + l)
        close(gd.nodes[l].data["energy"], ref["tuple_energy"][l], atol=1e-4, msg="tuple energy " + l)
    gscale = float(ref["gradient"].abs().max())
    close(gd.nodes["n1"].data["gradient"], ref["gradient"], atol=2e-5 * gscale, msg="gradient")


def test_energy_matches_reference_golden(golden_dir):
    gold = np.load(os.path.join(golden_dir, "cfg1_forward.npz"))
    g = S.cfg1_graph(n_confs=5, seed=0)
    params = {l: {k: torch.as_tensor(gold[f"{l}_{k}"]) for k in (("k", "eq") if l in ("n2", "n3") else ("k",))} for l in LEVELS}
    gd = g.to(DEV)
    write_params(gd, params, DEV)
    gd = Energy(terms=["bond", "angle", "proper", "improper"])(gd)
    close(gd.nodes["g"].data["energy"], torch.as_tensor(gold["g_energy"]), atol=2e-2, msg="energy vs reference golden")
    close(gd.nodes["n1"].data["gradient"], torch.as_tensor(gold["gradient"]), atol=5e-3, msg="gradient vs reference golden")
    for l in LEVELS:
        close(gd.nodes[l].data["x"], torch.as_tensor(gold[f"{l}_x"]), atol=2e-6, msg="x " + l)


def test_energy_vjp_matches_double_backward():
    g = batch([S.trimer_molecule(s, 6) for s in (0, 1, 2)])
    params = random_params(g, seed=5)
    gen = torch.Generator().manual_seed(9)
    gE = torch.randn(3, 6, generator=gen)
    gG = torch.randn(g.num_nodes("n1"), 6, 3, generator=gen)
    leaf = {l: {k: v.clone().requires_grad_(True) for k, v in d.items()} for l, d in params.items()}
    ref = oracle_energy(g, leaf, create_graph=True)
    ((ref["energy"] * gE).sum() + (ref["gradient"] * gG).sum()).backward()
    gd = g.to(DEV)
    dleaf = {l: {k: v.clone().to(DEV).requires_grad_(True) for k, v in d.items()} for l, d in params.items()}
    for l in LEVELS:
        for key, v in dleaf[l].items():
            gd.nodes[l].data[key] = v
    gd = Energy(terms=["bond", "angle", "proper", "improper"], materialize_tuple_terms=False)(gd)
    ((gd.nodes["g"].data["energy"] * gE.to(DEV)).sum() + (gd.nodes["n1"].data["gradient"] * gG.to(DEV)).sum()).backward()
    for l in LEVELS:
        for key in params[l]:
            r = leaf[l][key].grad
            close(dleaf[l][key].grad, r, atol=3e-5 * float(r.abs().max()), msg=f"d{key} {l}")


@pytest.mark.parametrize("n_parts", [1, 4, 7])
def test_energy_matches_oracle_at_sweep_shape(n_parts):
    """cfg-4 shape family (200 molecules x 500 conformations: 16 conformation chunks per molecule, TMA-staged rows)
    against the oracle, element-wise, for several program partitions (warps per CTA)."""
    g = S.cfg4_batch(n_mol=200, n_confs=500, seed=3)
    params = random_params(g, seed=4)
    ref = oracle_energy(g, params)
    gd = g.to(DEV)
    write_params(gd, params, DEV)
    gd._plan_cache = {"plan": build_plan(g, n_parts=n_parts).to(DEV)}
    out = Energy(materialize_tuple_terms=False)(gd)
    escale = float(ref["energy"].abs().max())
    close(out.nodes["g"].data["energy"], ref["energy"], atol=1e-5 * escale, msg="total energy")
    for l, term in zip(LEVELS, ("bond", "angle", "proper", "improper")):
        close(out.nodes["g"].data["energy_" + term], ref["term_energy"][l], atol=1e-5 * escale, msg="term energy " + term)
    gscale = float(ref["gradient"].abs().max())
    close(out.nodes["n1"].data["gradient"], ref["gradient"], atol=2e-5 * gscale, msg="gradient")


def test_energy_unaligned_conformation_counts_match_aligned():
    """C % 4 != 0 takes the coalesced-load path instead of the TMA rows: same numbers for the shared conformations."""
    g = S.cfg4_batch(n_mol=3, n_confs=52, seed=5)
    params = random_params(g, seed=6)
    outs = {}
    for C in (52, 50, 33, 1):
        gd = g.to(DEV)
        gd.nodes["n1"].data["xyz"] = gd.nodes["n1"].data["xyz"][:, :C].contiguous()
        write_params(gd, params, DEV)
        o = Energy(materialize_tuple_terms=False)(gd)
        outs[C] = (o.nodes["g"].data["energy"].clone(), o.nodes["n1"].data["gradient"].clone())
    for C in (50, 33, 1):
        assert torch.equal(outs[C][0], outs[52][0][:, :C]) and torch.equal(outs[C][1], outs[52][1][:, :C])


def test_energy_properties_full_size():
    """cfg 4 shape (scaled to 200 molecules x 500 conformations): determinism, translation invariance of energies,
    zero net force per molecule and conformation, linearity in the force constants."""
    g = S.cfg4_batch(n_mol=200, n_confs=500, seed=0)
    params = random_params(g, seed=1)
    gd = g.to(DEV)
    write_params(gd, params, DEV)
    en = Energy(materialize_tuple_terms=False)
    out = en(gd)
    e1, f1 = out.nodes["g"].data["energy"].clone(), out.nodes["n1"].data["gradient"].clone()
    out = en(gd)
    assert torch.equal(out.nodes["g"].data["energy"], e1) and torch.equal(out.nodes["n1"].data["gradient"], f1), "not deterministic"
    assert bool(torch.isfinite(e1).all()) and bool(torch.isfinite(f1).all())
    net = f1.view(200, 40, 500, 3).sum(1)
    assert float(net.abs().max()) < 2e-2 * float(f1.abs().mean()) * 40 ** 0.5 + 1e-2, "net force per molecule is not ~0"
    gd.nodes["n1"].data["xyz"] = gd.nodes["n1"].data["xyz"] + torch.tensor([0.3, -1.2, 2.0], device=DEV)
    e2 = en(gd).nodes["g"].data["energy"]
    close(e2, e1, rtol=2e-4, atol=2e-3 * float(e1.abs().mean()), msg="translation invariance")
    gd.nodes["n1"].data["xyz"] = g.nodes["n1"].data["xyz"].to(DEV)
    for l in LEVELS:
        gd.nodes[l].data["k"] = gd.nodes[l].data["k"] * 2
    e3 = en(gd).nodes["g"].data["energy"]
    close(e3, 2 * e1, rtol=1e-5, atol=1e-3, msg="linearity in k")


def test_energy_api_errors_and_term_subsets():
    g = S.cfg1_graph(n_confs=3)
    params = random_params(g)
    gd = g.to(DEV)
    with pytest.raises(RuntimeError, match="has no k"):
        Energy()(gd)
    write_params(gd, params, DEV)
    xyz = gd.nodes["n1"].data.pop("xyz")
    with pytest.raises(ValueError, match="xyz"):
        Energy()(gd)
    gd.nodes["n1"].data["xyz"] = xyz
    with pytest.raises(ValueError, match="terms must be a list"):
        Energy(terms="bond")
    # subsets and per-term gradient contributions add up to the total
    full = Energy()(gd)
    tot_e, tot_g = full.nodes["g"].data["energy"].clone(), full.nodes["n1"].data["gradient"].clone()
    sub = Energy(terms=["bond", "angle"], write_suffix="_ba")(gd)
    close(sub.nodes["g"].data["energy_ba"], full.nodes["g"].data["energy_bond"] + full.nodes["g"].data["energy_angle"], atol=1e-3)
    contrib = Energy(gradient_contributions=True)(gd)
    s = sum(contrib.nodes["n1"].data["gradient_" + t] for t in ("bond", "angle", "proper", "improper"))
    close(s, tot_g, rtol=1e-4, atol=1e-3, msg="gradient contributions")
    with torch.no_grad():                      # usable under no_grad, still writes gradients (energy.py:118-120)
        ng = Energy()(gd)
    assert torch.equal(ng.nodes["n1"].data["gradient"], tot_g) and torch.equal(ng.nodes["g"].data["energy"], tot_e)
