"""
TEST INFRASTRUCTURE.  Generates tests/golden/*.npz|json by running the UNMODIFIED reference
(/root/reference/src, on oracle/dgl_standin) in the build container:

    python oracle/make_golden.py

Fixtures:
  state_dict_layout.json   key -> [shape, dtype] of GrappaModel(**configs/model/default.yaml).state_dict()
  state_dict_constants.npz the non-learnable buffers of that state dict (statistics, permutations, encodings)
  cfg1_forward.npz         22-atom ACE-ALA-NME, 5 confs: inputs, h, k/eq per level, internal coordinates,
                           per-tuple / per-term / total energies, dE/dxyz  (dihedral noise zeroed)
  batch3_train.npz         3 x 40-atom molecules x 4 confs batched with the reference's dgl_utils.batch:
                           batched idxs/offsets, MolwiseLoss value, and per-parameter gradient digests after backward
  tuples.npz               angle / proper / improper rows from utils/tuple_indices.py for four topologies
Weights are oracle.weights.deterministic_fill_ applied to the reference model (no weight files are stored).
"""
import contextlib
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import reference_loader as rl  # noqa: E402
from weights import deterministic_fill_  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


@contextlib.contextmanager
def no_dihedral_noise():
    """internal_coordinates.py:178-180 adds randn*1e-5; zero it for the deterministic golden."""
    orig = torch.randn_like
    torch.randn_like = lambda x, *a, **k: torch.zeros_like(x)
    try:
        yield
    finally:
        torch.randn_like = orig


def to_reference_graph(hg):
    """grappa_b200.graph.HeteroGraph (single molecule) -> stand-in DGL heterograph with Molecule.to_dgl's layout."""
    import dgl
    src, dst = hg.edges()
    data = {("n1", "n1_edge", "n1"): (src, dst)}
    for t in ["n2", "n3", "n4", "n4_improper", "g"]:
        n = hg.num_nodes(t)
        data[(t, f"{t}_edge", t)] = (torch.arange(n), torch.arange(n))
    g = dgl.heterograph(data, num_nodes_dict={t: hg.num_nodes(t) for t in hg.ntypes})
    for t in hg.ntypes:
        for k, v in hg.nodes[t].data.items():
            g.nodes[t].data[k] = v.clone()
    return g


def digest(t: torch.Tensor):
    f = t.detach().double().flatten()
    return np.array([f.sum().item(), f.abs().sum().item(), (f * torch.arange(1, len(f) + 1, dtype=torch.float64) % 7.0).sum().item()]
                    + f[:8].tolist() + [0.0] * max(0, 8 - len(f)))


def main():
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(8)
    rl.load_reference()
    from grappa.models.grappa import GrappaModel
    from grappa.models.energy import Energy
    from grappa.training.loss import MolwiseLoss
    from grappa.utils import dgl_utils, tuple_indices
    from grappa_b200 import synthetic as S

    # importing grappa.training runs experiment.py:7 `torch.set_float32_matmul_precision('medium')`, which lets oneDNN
    # use bf16 inside fp32 matmuls on this CPU (errors ~1e-3).  The oracle setting is full fp32.
    torch.set_float32_matmul_precision("highest")

    cfg = rl.default_model_config()
    torch.manual_seed(0)
    model = GrappaModel(**cfg).eval()
    sd = model.state_dict()
    with open(os.path.join(OUT, "state_dict_layout.json"), "w") as f:
        json.dump({k: [list(v.shape), str(v.dtype)] for k, v in sd.items()}, f, indent=0)
    # the buffers that deterministic_fill_ leaves alone (parameter statistics, permutations, positional encodings ...):
    # with them and the layout a state dict can be rebuilt anywhere without the reference or the product's model code
    from weights import _KEEP
    np.savez_compressed(os.path.join(OUT, "state_dict_constants.npz"),
                        **{k: v.numpy() for k, v in sd.items() if not torch.is_floating_point(v) or any(c in k for c in _KEEP)})
    if "--constants-only" in sys.argv:
        return
    deterministic_fill_(sd)
    model.load_state_dict(sd)

    # ---- cfg 1 forward ----
    hg = S.cfg1_graph(n_confs=5, seed=0)
    g = to_reference_graph(hg)
    with no_dihedral_noise():
        g = torch.nn.Sequential(model, Energy(suffix="", terms=["bond", "angle", "proper", "improper"]))(g)
    out = {}
    for k, v in hg.nodes["n1"].data.items():
        out["in_n1_" + k] = v.numpy()
    src, dst = hg.edges()
    out["in_edges"] = torch.stack([src, dst]).numpy()
    for lvl in ["n2", "n3", "n4", "n4_improper"]:
        out[f"in_{lvl}_idxs"] = hg.nodes[lvl].data["idxs"].numpy()
        d = g.nodes[lvl].data
        out[f"{lvl}_k"] = d["k"].detach().numpy()
        if "eq" in d:
            out[f"{lvl}_eq"] = d["eq"].detach().numpy()
        out[f"{lvl}_x"] = d["x"].detach().numpy()
        out[f"{lvl}_energy"] = d["energy"].detach().numpy()
    out["h"] = g.nodes["n1"].data["h"].detach().numpy()
    out["gradient"] = g.nodes["n1"].data["gradient"].detach().numpy()
    for k, v in g.nodes["g"].data.items():
        if k.startswith("energy"):
            out["g_" + k] = v.detach().numpy()
    np.savez_compressed(os.path.join(OUT, "cfg1_forward.npz"), **out)
    print("cfg1: energy", out["g_energy"][0, :3], "k2", out["n2_k"][:3], "|grad|", np.abs(out["gradient"]).mean())

    # ---- 3-molecule training batch ----
    model.train()  # dropout must be off for a deterministic golden: set p=0 on every Dropout module
    for m in model.modules():
        if isinstance(m, torch.nn.Dropout):
            m.p = 0.0
    mols = [S.trimer_molecule(seed=s, n_confs=4) for s in (0, 1, 2)]
    refg = [to_reference_graph(m) for m in mols]
    bg = dgl_utils.batch(refg)
    out = {}
    for lvl in ["n2", "n3", "n4", "n4_improper"]:
        out[f"batched_{lvl}_idxs"] = bg.nodes[lvl].data["idxs"].numpy()
        out[f"batch_num_{lvl}"] = bg.batch_num_nodes(lvl).numpy()
    out["batch_num_n1"] = bg.batch_num_nodes("n1").numpy()
    e_src, e_dst = bg.edges(etype="n1_edge")
    out["batched_edges"] = torch.stack([e_src, e_dst]).numpy()
    with no_dihedral_noise():
        bg = torch.nn.Sequential(model, Energy(suffix="", terms=["bond", "angle", "proper", "improper"]))(bg)
    loss_fn = MolwiseLoss(gradient_weight=0.8, energy_weight=1.0, param_weight=0.0, proper_regularisation=1.0,
                          improper_regularisation=1.0)
    loss = loss_fn(bg)
    loss.backward()
    out["loss"] = np.array(loss.item())
    out["energy"] = bg.nodes["g"].data["energy"].detach().numpy()
    out["gradient"] = bg.nodes["n1"].data["gradient"].detach().numpy()
    names, digs = [], []
    for k, p in model.named_parameters():
        if p.grad is None:
            continue
        names.append(k)
        digs.append(digest(p.grad))
    out["grad_names"] = np.array(names)
    out["grad_digests"] = np.stack(digs)
    none_grads = [k for k, p in model.named_parameters() if p.grad is None]
    out["grad_none"] = np.array(none_grads)
    np.savez_compressed(os.path.join(OUT, "batch3_train.npz"), **out)
    print("batch3: loss", loss.item(), "n grads", len(names), "None:", none_grads)

    # ---- tuple enumeration ----
    out = {}
    rng = np.random.default_rng(1)
    tmpls = {"ace": S.template_ace_ala_nme(), "AFA": S.template_polymer("AFA"), "poly20": S.template_polymer(S.alternating_sequence(20))}
    sh = S.template_polymer("FAF")
    sh["bonds"] = sh["bonds"][rng.permutation(len(sh["bonds"]))][:, ::-1].copy()
    tmpls["shuffledFAF"] = sh
    for name, t in tmpls.items():
        bonds = [tuple(b) for b in t["bonds"].tolist()]
        nd = tuple_indices.get_neighbor_dict(bonds, sort=True)
        r = tuple_indices.get_idx_tuples(bonds, neighbor_dict=nd, is_sorted=True)
        nb = S._neighbour_lists(t["bonds"], len(t["atomic_numbers"]))
        imp_in = [(nb[c][0], nb[c][1], c, nb[c][2]) for c in t["improper_centres"].tolist()]
        _, imp = tuple_indices.get_torsions(imp_in, nd)
        out[name + "_bonds"] = t["bonds"]
        out[name + "_centres"] = t["improper_centres"]
        out[name + "_angles"] = np.array(r["angles"], np.int64).reshape(-1, 3)
        out[name + "_propers"] = np.array(r["propers"], np.int64).reshape(-1, 4)
        out[name + "_impropers"] = np.array(imp, np.int64).reshape(-1, 4)
    np.savez_compressed(os.path.join(OUT, "tuples.npz"), **out)
    print("golden written to", OUT)


if __name__ == "__main__":
    main()
