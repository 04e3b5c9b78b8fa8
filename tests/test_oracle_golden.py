"""Pins the CPU oracle (oracle/grappa_oracle.py) to golden vectors produced by the unmodified reference
(oracle/make_golden.py).  CPU only."""
import json
import os

import numpy as np
import pytest
import torch

import grappa_oracle as O
from weights import deterministic_fill_
from grappa_b200 import synthetic as S
from grappa_b200.graph import batch

RTOL = 1e-4


def _sd_from_layout(golden_dir):
    with open(os.path.join(golden_dir, "state_dict_layout.json")) as f:
        layout = json.load(f)
    from grappa_b200.model import GrappaModel
    m = GrappaModel.from_default_config()
    sd = m.state_dict()
    assert list(sd.keys()) == list(layout.keys())
    deterministic_fill_(sd)
    return {k: v.clone() for k, v in sd.items()}


def test_cfg1_forward_matches_reference(golden_dir):
    gold = np.load(os.path.join(golden_dir, "cfg1_forward.npz"))
    sd = _sd_from_layout(golden_dir)
    g = S.cfg1_graph(n_confs=5, seed=0)
    # synthetic inputs are regenerated, and must equal what the golden run saw
    for k, v in g.nodes["n1"].data.items():
        np.testing.assert_array_equal(v.numpy(), gold["in_n1_" + k])
    for lvl in O.LEVELS:
        np.testing.assert_array_equal(g.nodes[lvl].data["idxs"].numpy(), gold[f"in_{lvl}_idxs"])
    with torch.no_grad():
        h, params = O.model_forward(sd, g)
    np.testing.assert_allclose(h.numpy(), gold["h"], rtol=RTOL, atol=1e-5)
    for lvl in O.LEVELS:
        np.testing.assert_allclose(params[lvl]["k"].numpy(), gold[f"{lvl}_k"], rtol=RTOL, atol=1e-5)
        if "eq" in params[lvl]:
            np.testing.assert_allclose(params[lvl]["eq"].numpy(), gold[f"{lvl}_eq"], rtol=RTOL, atol=1e-6)
    n1, src, dst, idxs, counts = O.graph_inputs(g)
    en = O.energy_forward(n1["xyz"], idxs, params, counts)
    for lvl in O.LEVELS:
        np.testing.assert_allclose(en["x"][lvl].detach().numpy(), gold[f"{lvl}_x"], rtol=RTOL, atol=1e-6)
        np.testing.assert_allclose(en["tuple_energy"][lvl].detach().numpy(), gold[f"{lvl}_energy"], rtol=RTOL, atol=1e-4)
    for lvl, term in zip(O.LEVELS, ("bond", "angle", "proper", "improper")):
        np.testing.assert_allclose(en["term_energy"][lvl].numpy(), gold[f"g_energy_{term}"], rtol=RTOL, atol=1e-3)
    np.testing.assert_allclose(en["energy"].detach().numpy(), gold["g_energy"], rtol=RTOL, atol=1e-3)
    np.testing.assert_allclose(en["gradient"].detach().numpy(), gold["gradient"], rtol=RTOL, atol=2e-3)


def _digest(t):
    f = t.detach().double().flatten()
    return np.array([f.sum().item(), f.abs().sum().item(),
                     (f * torch.arange(1, len(f) + 1, dtype=torch.float64) % 7.0).sum().item()]
                    + f[:8].tolist() + [0.0] * max(0, 8 - len(f)))


def test_batch3_loss_and_grads_match_reference(golden_dir):
    gold = np.load(os.path.join(golden_dir, "batch3_train.npz"))
    sd = _sd_from_layout(golden_dir)
    params = {k: v.requires_grad_(True) for k, v in sd.items() if torch.is_floating_point(v)
              and not any(s in k for s in ("to_k.", "to_eq.", "k_mean", "k_std", "prefactors", "positional"))}
    sd.update(params)
    g = batch([S.trimer_molecule(seed=s, n_confs=4) for s in (0, 1, 2)])
    for lvl in O.LEVELS:
        np.testing.assert_array_equal(g.nodes[lvl].data["idxs"].numpy(), gold[f"batched_{lvl}_idxs"])
        np.testing.assert_array_equal(g.batch_num_nodes(lvl).numpy(), gold[f"batch_num_{lvl}"])
    src, dst = g.edges()
    np.testing.assert_array_equal(torch.stack([src, dst]).numpy(), gold["batched_edges"])
    loss, prm, en = O.train_step_loss(sd, g)
    np.testing.assert_allclose(loss.item(), gold["loss"], rtol=RTOL)
    np.testing.assert_allclose(en["energy"].detach().numpy(), gold["energy"], rtol=RTOL, atol=1e-2)
    np.testing.assert_allclose(en["gradient"].detach().numpy(), gold["gradient"], rtol=RTOL, atol=5e-3)
    loss.backward()
    names = [str(n) for n in gold["grad_names"]]
    seen = set()
    for name, dg in zip(names, gold["grad_digests"]):
        if name.startswith("gnn.blocks."):  # alias of gnn.att_blocks.*
            continue
        seen.add(name)
        mine = _digest(sd[name].grad)
        scale = max(dg[1] / max(sd[name].numel(), 1), 1e-6)  # mean |grad|
        np.testing.assert_allclose(mine[:2], dg[:2], rtol=2e-3, atol=scale * sd[name].numel() * 2e-4, err_msg=name)
        np.testing.assert_allclose(mine[3:], dg[3:], rtol=2e-3, atol=scale * 2e-2, err_msg=name)
    for name in gold["grad_none"]:
        assert sd[str(name)].grad is None


@pytest.mark.reference
def test_python_default_config_matches_live_reference():
    """GrappaModel() with the Python defaults (ungated torsions, n_periodicity_improper = 3: models/grappa.py:28-29)
    against the live reference; skipped where /root/reference is absent."""
    import reference_loader as rl
    if not rl.reference_available():
        pytest.skip("reference tree not present")
    rl.load_reference()
    from grappa.models.grappa import GrappaModel as RefModel
    from make_golden import to_reference_graph
    from grappa_b200.model import GrappaModel
    torch.set_float32_matmul_precision("highest")
    ref = RefModel().eval()
    mine = GrappaModel()
    sd = ref.state_dict()
    assert list(sd.keys()) == list(mine.state_dict().keys())
    deterministic_fill_(sd)
    ref.load_state_dict(sd)
    hg = S.cfg1_graph(n_confs=2, seed=3)
    with torch.no_grad():
        g = ref(to_reference_graph(hg))
        h, params = O.model_forward({k: v.clone() for k, v in sd.items()}, hg)
    for lvl in O.LEVELS:
        for key, v in params[lvl].items():
            np.testing.assert_allclose(v.numpy(), g.nodes[lvl].data[key].numpy(), rtol=1e-5, atol=1e-6, err_msg=f"{lvl} {key}")
    assert params["n4_improper"]["k"].shape == (12, 3)
