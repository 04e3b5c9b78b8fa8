"""End-to-end parity of the drop-in modules: GrappaModel / Energy / fused loss against the reference's golden vectors
(tests/golden, produced by the unmodified reference) and against the CPU oracle on larger seeded inputs."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import grappa_oracle as O
from weights import deterministic_fill_
from grappa_b200 import ops, synthetic as S
from grappa_b200.energy import Energy
from grappa_b200.graph import batch
from grappa_b200.model import GrappaModel
from grappa_b200.plan import get_plan

DEV = "cuda"
LEVELS = ("n2", "n3", "n4", "n4_improper")
K_STD_TORSION = 1.3   # scale of gated torsion outputs (k_std ~ 0.25..1.35): atol for torsion k is 1e-4 * k_std


def build_model(train=False):
    m = GrappaModel.from_default_config()
    sd = m.state_dict()
    deterministic_fill_(sd)
    m.load_state_dict(sd, strict=True)
    m = m.to(DEV)
    return (m.train() if train else m.eval()), {k: v.clone() for k, v in sd.items()}


def assert_params_close(got, want, tag=""):
    for lvl in LEVELS:
        for key, ref in want[lvl].items():
            a = got[lvl][key].detach().cpu().numpy()
            b = ref.detach().cpu().numpy() if isinstance(ref, torch.Tensor) else ref
            if lvl in ("n4", "n4_improper"):
                # entries that land on opposite sides of the hard cutoff (|k| <= 1e-4 -> 0) are compared with atol only
                np.testing.assert_allclose(a, b, rtol=1e-4, atol=2e-4 * K_STD_TORSION, err_msg=f"{tag} {lvl} {key}")
            else:
                np.testing.assert_allclose(a, b, rtol=1e-4, atol=1e-6, err_msg=f"{tag} {lvl} {key}")


def graph_params(g):
    return {l: {k: g.nodes[l].data[k] for k in (("k", "eq") if l in ("n2", "n3") else ("k",))} for l in LEVELS}


def test_cfg1_forward_matches_reference_golden(golden_dir):
    gold = np.load(os.path.join(golden_dir, "cfg1_forward.npz"))
    model, _ = build_model()
    g = S.cfg1_graph(n_confs=5, seed=0).to(DEV)
    with torch.no_grad():
        g = torch.nn.Sequential(model, Energy(terms=["bond", "angle", "proper", "improper"]))(g)
    np.testing.assert_allclose(g.nodes["n1"].data["h"].cpu().numpy(), gold["h"], rtol=1e-4, atol=2e-5, err_msg="atom embedding h")
    want = {l: {k: gold[f"{l}_{k}"] for k in (("k", "eq") if l in ("n2", "n3") else ("k",))} for l in LEVELS}
    assert_params_close(graph_params(g), want, "golden")
    np.testing.assert_allclose(g.nodes["g"].data["energy"].cpu().numpy(), gold["g_energy"], rtol=1e-4, atol=5e-2)
    np.testing.assert_allclose(g.nodes["n1"].data["gradient"].cpu().numpy(), gold["gradient"], rtol=1e-4, atol=2e-2)
    assert g.nodes["n4"].data["k"].shape == (41, 3) and g.nodes["n4_improper"].data["k"].shape == (12, 2)
    assert float(g.nodes["n4_improper"].data["k"][:, 0].abs().max()) == 0.0     # only n = 2 is predicted


def test_batched_forward_matches_oracle_and_is_batch_invariant():
    model, sd = build_model()
    mols = [S.trimer_molecule(s, 3) for s in range(5)]
    g = batch(mols)
    with torch.no_grad():
        h_ref, p_ref = O.model_forward(sd, g)
        gd = model(g.to(DEV))
    np.testing.assert_allclose(gd.nodes["n1"].data["h"].cpu().numpy(), h_ref.numpy(), rtol=1e-4, atol=2e-5)
    assert_params_close(graph_params(gd), p_ref, "batched")
    # batch / unbatch invariance (what the reference's tests/test_dataloader.py:34-50 pins)
    with torch.no_grad():
        single = model(mols[2].to(DEV))
    lo = sum(m.num_nodes("n2") for m in mols[:2])
    np.testing.assert_allclose(single.nodes["n2"].data["k"].cpu().numpy(),
                               gd.nodes["n2"].data["k"][lo:lo + mols[2].num_nodes("n2")].cpu().numpy(), rtol=2e-5, atol=1e-4)


def _digest(t):
    f = t.detach().double().flatten().cpu()
    return np.array([f.sum().item(), f.abs().sum().item(),
                     (f * torch.arange(1, len(f) + 1, dtype=torch.float64) % 7.0).sum().item()]
                    + f[:8].tolist() + [0.0] * max(0, 8 - len(f)))


def train_loss(model, g, plan_device=DEV):
    g = torch.nn.Sequential(model, Energy(terms=["bond", "angle", "proper", "improper"], materialize_tuple_terms=False))(g)
    plan = get_plan(g)
    loss = ops.MolwiseLossFn.apply(g.nodes["g"].data["energy"], g.nodes["n1"].data["gradient"], g.nodes["n4"].data["k"],
                                   g.nodes["n4_improper"].data["k"], g.nodes["g"].data["energy_ref"],
                                   g.nodes["n1"].data["gradient_ref"], g.nodes["g"].data.get("is_dummy"), plan,
                                   (1.0, 0.8, 1.0, 1.0))
    return loss, g


def test_training_step_matches_reference_golden(golden_dir):
    """Loss value and all 216 parameter gradients of a 3-molecule batch vs the reference's MolwiseLoss.backward()."""
    gold = np.load(os.path.join(golden_dir, "batch3_train.npz"))
    model, _ = build_model(train=True)
    for m in model.modules():
        if isinstance(m, torch.nn.Dropout):
            m.p = 0.0
    g = batch([S.trimer_molecule(seed=s, n_confs=4) for s in (0, 1, 2)]).to(DEV)
    loss, g = train_loss(model, g)
    np.testing.assert_allclose(loss.item(), float(gold["loss"]), rtol=1e-4)
    np.testing.assert_allclose(g.nodes["g"].data["energy"].detach().cpu().numpy(), gold["energy"], rtol=1e-4, atol=5e-2)
    loss.backward()
    named = dict(model.named_parameters())
    worst = []
    for name, dg in zip([str(n) for n in gold["grad_names"]], gold["grad_digests"]):
        if name.startswith("gnn.blocks."):
            continue
        p = named[name]
        assert p.grad is not None, name
        mine = _digest(p.grad)
        mean_abs = max(dg[1] / p.numel(), 1e-9)
        e1 = abs(mine[1] - dg[1]) / max(dg[1], 1e-9)                     # sum |g|
        e2 = np.abs(mine[3:] - dg[3:]).max() / max(np.abs(dg[3:]).max(), mean_abs)
        worst.append((max(e1, e2), name))
        assert e1 < 1e-3 and e2 < 2e-3, f"{name}: rel err sum|g| {e1:.2e}, first entries {e2:.2e}"
    for name in gold["grad_none"]:
        assert named[str(name)].grad is None, name
    print("worst gradient digests:", sorted(worst)[-3:])


def test_training_step_gradients_match_oracle_elementwise():
    model, sd = build_model(train=True)
    for m in model.modules():
        if isinstance(m, torch.nn.Dropout):
            m.p = 0.0
    g = batch([S.trimer_molecule(seed=s, n_confs=3) for s in (5, 6)])
    leaf = {k: (v.requires_grad_(True) if torch.is_floating_point(v) and not any(s in k for s in ("to_k.", "to_eq.", "k_mean", "k_std", "prefactors", "positional")) else v)
            for k, v in sd.items()}
    loss_ref, _, _ = O.train_step_loss(leaf, g)
    loss_ref.backward()
    loss, _ = train_loss(model, g.to(DEV))
    loss.backward()
    np.testing.assert_allclose(loss.item(), loss_ref.item(), rtol=1e-4)
    for name, p in model.named_parameters():
        r = leaf[name].grad
        if r is None:
            assert p.grad is None, name
            continue
        scale = float(r.abs().max())
        np.testing.assert_allclose(p.grad.cpu().numpy(), r.numpy(), rtol=2e-3, atol=2e-4 * scale, err_msg=name)


def test_training_step_at_the_benchmarked_size_matches_oracle():
    """BASELINE.json configs[1] exactly as bench.py times it (32 molecules x 40 atoms x 50 conformations), dropout off:
    loss, energies, forces at rtol 1e-4 and every parameter gradient against the CPU oracle."""
    model, sd = build_model(train=True)
    for m in model.modules():
        if isinstance(m, torch.nn.Dropout):
            m.p = 0.0
    g = S.cfg2_batch(32, 50)
    leaf = {k: (v.requires_grad_(True) if torch.is_floating_point(v) and not any(s in k for s in ("to_k.", "to_eq.", "k_mean", "k_std", "prefactors", "positional")) else v)
            for k, v in sd.items()}
    loss_ref, params_ref, en_ref = O.train_step_loss(leaf, g)
    loss_ref.backward()
    loss, gd = train_loss(model, g.to(DEV))
    loss.backward()
    np.testing.assert_allclose(loss.item(), loss_ref.item(), rtol=1e-4)
    assert_params_close(graph_params(gd), params_ref, "cfg2")
    e, er = gd.nodes["g"].data["energy"].detach().cpu().numpy(), en_ref["energy"].detach().numpy()
    np.testing.assert_allclose(e, er, rtol=1e-4, atol=1e-5 * np.abs(er).max())
    f, fr = gd.nodes["n1"].data["gradient"].detach().cpu().numpy(), en_ref["gradient"].detach().numpy()
    np.testing.assert_allclose(f, fr, rtol=1e-4, atol=2e-5 * np.abs(fr).max())
    for name, p in model.named_parameters():
        r = leaf[name].grad
        if r is None:
            assert p.grad is None, name
            continue
        scale = float(r.abs().max())
        np.testing.assert_allclose(p.grad.cpu().numpy(), r.numpy(), rtol=2e-3, atol=2e-4 * scale, err_msg=name)


def test_skipped_writer_does_not_reapply_stale_gradients():
    """A batch without impropers must not leave the previous step's improper-writer gradients in the flat buffer (they
    would be all-reduced and applied again): step with impropers, then without, for eager and captured steps."""
    from grappa_b200.batching import PackedBatch
    from grappa_b200.train_step import TrainStep

    def without_impropers(seed):
        g = S.trimer_molecule(seed=seed, n_confs=3)
        g.nodes["n4_improper"].data["idxs"] = torch.zeros((0, 4), dtype=torch.long)
        g._num_nodes["n4_improper"] = 0
        g._batch_num_nodes["n4_improper"] = torch.tensor([0])
        return g

    full = PackedBatch(batch([S.trimer_molecule(seed=s, n_confs=3) for s in (0, 1)]))
    bare = PackedBatch(batch([without_impropers(s) for s in (0, 1)]))
    for use_graph in (False, True):
        model, _ = build_model(train=True)
        ts = TrainStep(model, DEV, use_cuda_graph=use_graph, improper_regularisation=0.0)
        lo, hi = ts._writer_ranges["n4_improper"]
        for _ in range(5 if use_graph else 1):
            ts.step(full)
        assert float(ts.flat.grads[lo:hi].abs().max()) > 0
        before = ts.flat.params[lo:hi].clone()
        for _ in range(5 if use_graph else 1):
            ts.step(bare)
            assert float(ts.flat.grads[lo:hi].abs().max()) == 0.0, "stale improper-writer gradients survived"
        ts.synchronize()
        # with zero gradients the parameters only coast on Adam's momentum: bounded by lr per step
        assert float((ts.flat.params[lo:hi] - before).abs().max()) <= 5 * ts.lr * 1.01


def test_dropout_training_mode_runs_and_differs():
    model, _ = build_model(train=True)
    g = batch([S.trimer_molecule(seed=s, n_confs=2) for s in (0, 1)]).to(DEV)
    ops.DropoutSeeds.manual_seed(1)
    k1 = model(g).nodes["n2"].data["k"].detach().clone()
    k2 = model(g).nodes["n2"].data["k"].detach().clone()
    assert bool(torch.isfinite(k1).all()) and not torch.equal(k1, k2)
    ops.DropoutSeeds.manual_seed(1)
    k3 = model(g).nodes["n2"].data["k"].detach()
    assert torch.equal(k1, k3), "dropout must be reproducible for a fixed seed"
    loss, _ = train_loss(model, g)
    loss.backward()
    assert all(bool(torch.isfinite(p.grad).all()) for p in model.parameters() if p.grad is not None)


def test_api_errors_and_edge_cases():
    model, _ = build_model()
    g = S.cfg1_graph(n_confs=1)
    g.nodes["n2"].data["idxs"] = g.nodes["n2"].data["idxs"].clone()
    g.nodes["n2"].data["idxs"][0, 1] = 99
    with pytest.raises(AssertionError, match="Encountered idxs up to 99"):
        model(g.to(DEV))
    # a graph without impropers / angles types is skipped like the reference does (interaction_parameters.py:374,572)
    g = S.cfg1_graph(n_confs=1)
    for t in ("n3", "n4_improper"):
        del g._num_nodes[t], g._ndata[t], g._batch_num_nodes[t]
    with torch.no_grad():
        out = model(g.to(DEV))
    assert "k" in out.nodes["n2"].data and "k" in out.nodes["n4"].data
    # zero-length tuple sets give empty outputs (:179-180, :579-580)
    g = S.cfg1_graph(n_confs=1)
    g.nodes["n4"].data["idxs"] = torch.zeros((0, 4), dtype=torch.long)
    g._num_nodes["n4"] = 0
    g._batch_num_nodes["n4"] = torch.tensor([0])
    with torch.no_grad():
        out = model(g.to(DEV))
    assert out.nodes["n4"].data["k"].shape == (0, 3)
    with pytest.raises(Exception):
        model(S.cfg1_graph(n_confs=1))          # CPU graph: no fallback


def test_cfg3_inference_matches_oracle_at_scale():
    """BASELINE cfg 3 shape (scaled to 200 residues = 3000 atoms so that the CPU oracle stays within seconds):
    parameters of a long polymer, one graph, no_grad — what Grappa.predict runs."""
    model, sd = build_model()
    g = S.cfg3_graph(n_res=200, seed=1)
    with torch.no_grad():
        h_ref, p_ref = O.model_forward(sd, g)
        gd = model(g.to(DEV))
    np.testing.assert_allclose(gd.nodes["n1"].data["h"].cpu().numpy(), h_ref.numpy(), rtol=1e-4, atol=3e-5)
    assert_params_close(graph_params(gd), p_ref, "cfg3")
    assert gd.nodes["n4"].data["k"].shape == (g.num_nodes("n4"), 3)


def test_cfg3_full_size_properties():
    """Full 15,000-atom system: finite outputs, physical ranges of the range maps, permutation symmetry of the readout
    (reversing a bond / angle / proper tuple leaves its parameters unchanged), determinism."""
    model, _ = build_model()
    g = S.cfg3_graph()
    gd = g.to(DEV)
    with torch.no_grad():
        out = model(gd)
    k2, eq2 = out.nodes["n2"].data["k"].clone(), out.nodes["n2"].data["eq"].clone()
    k3, eq3 = out.nodes["n3"].data["k"].clone(), out.nodes["n3"].data["eq"].clone()
    k4 = out.nodes["n4"].data["k"].clone()
    for t in (k2, eq2, k3, eq3, k4, out.nodes["n4_improper"].data["k"]):
        assert bool(torch.isfinite(t).all())
    assert float(k2.min()) > 0 and float(eq2.min()) > 0 and float(k3.min()) > 0
    assert 0 < float(eq3.min()) and float(eq3.max()) < 3.1416
    g2 = S.cfg3_graph()
    for lvl in ("n2", "n3", "n4"):
        g2.nodes[lvl].data["idxs"] = g2.nodes[lvl].data["idxs"].flip(1).contiguous()
    with torch.no_grad():
        out2 = model(g2.to(DEV))
    np.testing.assert_allclose(out2.nodes["n2"].data["k"].cpu().numpy(), k2.cpu().numpy(), rtol=2e-5, atol=1e-4)
    np.testing.assert_allclose(out2.nodes["n3"].data["eq"].cpu().numpy(), eq3.cpu().numpy(), rtol=2e-5, atol=1e-6)
    np.testing.assert_allclose(out2.nodes["n4"].data["k"].cpu().numpy(), k4.cpu().numpy(), rtol=1e-4, atol=2e-4 * K_STD_TORSION)
    with torch.no_grad():
        out3 = model(S.cfg3_graph().to(DEV))
    assert torch.equal(out3.nodes["n2"].data["k"], k2) and torch.equal(out3.nodes["n4"].data["k"], k4), "not deterministic"


def test_python_default_config_ungated_torsions():
    """GrappaModel() Python defaults differ from the YAML (models/grappa.py:28-29): ungated torsions with the mean shift
    and n_periodicity_improper = 3 (only n = 2 predicted)."""
    m = GrappaModel()
    sd = m.state_dict()
    deterministic_fill_(sd)
    m.load_state_dict(sd)
    m = m.to(DEV).eval()
    g = batch([S.trimer_molecule(s, 2) for s in (0, 1)])
    with torch.no_grad():
        h_ref, p_ref = O.model_forward({k: v.clone() for k, v in sd.items()}, g)
        gd = m(g.to(DEV))
    assert gd.nodes["n4_improper"].data["k"].shape == (48, 3)
    assert float(gd.nodes["n4_improper"].data["k"][:, [0, 2]].abs().max()) == 0.0
    assert_params_close(graph_params(gd), p_ref, "python-default config")


@pytest.mark.parametrize("use_graph", [False, True])
def test_bucketed_gradient_exchange_choreography(use_graph):
    """TrainStep(overlap_comm=True) issues the writers' gradient bucket from an autograd boundary node at the atom
    embedding (on a communication stream, while the GNN backward runs) and the GNN bucket after backward.
    One GPU cannot run NCCL against a peer, so the exchange is replaced by a stand-in that doubles the bucket in place
    (what a 2-rank sum of identical gradients would give): every gradient must come out exactly doubled, every element
    exactly once, eagerly and through the captured graph, and with the exchange disabled the step is unchanged."""
    from grappa_b200.batching import PackedBatch
    from grappa_b200.train_step import TrainStep
    packed = PackedBatch(batch([S.trimer_molecule(seed=s, n_confs=5) for s in range(4)]))
    outs = {}
    for mode in ("plain", "overlap_noop", "overlap_double"):
        model, _ = build_model(train=True)
        ops.DropoutSeeds.manual_seed(1234)           # dropout stays ON: all runs must draw the same masks
        ts = TrainStep(model, DEV, use_cuda_graph=use_graph, overlap_comm=(mode != "plain"))
        assert 0 < ts.bucket_split < ts.flat.numel
        calls = []
        if mode == "overlap_double":
            def fake(view):
                calls.append((view.data_ptr(), view.numel()))
                view.mul_(2.0)
            ts._reduce_fn = fake
        for _ in range(5):                            # with graphs: 3 eager warm-ups, capture, replay
            loss = ts.compute_gradients(packed)
        torch.cuda.synchronize()
        outs[mode] = (float(loss), ts.flat.grads.clone())
        if mode == "overlap_double":
            assert len(calls) >= 2 and sum(n for _, n in calls[:2]) == ts.flat.numel     # two buckets tile the buffer
            assert calls[0][1] == ts.flat.numel - ts.bucket_split                          # writers first
    assert outs["plain"][0] == outs["overlap_noop"][0] == outs["overlap_double"][0]
    assert torch.equal(outs["plain"][1], outs["overlap_noop"][1])
    assert torch.equal(outs["plain"][1] * 2.0, outs["overlap_double"][1])


def test_token_deduplication_is_exact_in_forward_and_consistent_in_backward(monkeypatch):
    """Running the first LayerNorm + QKV projection once per distinct (atom, pe) row and gathering per token gives
    bit-identical parameters (row-wise arithmetic) and the same parameter gradients up to summation order."""
    import grappa_b200.model as M
    g_host = batch([S.trimer_molecule(seed=s, n_confs=3) for s in range(3)])
    results = {}
    for dedup in (False, True):
        monkeypatch.setattr(M, "DEDUP_TOKENS", dedup)
        model, _ = build_model(train=True)
        for m in model.modules():
            if isinstance(m, torch.nn.Dropout):
                m.p = 0.0
        calls = []            # the table path shows up as attention calls that gather from a table (no row_gather is left)
        orig = ops.tuple_attention
        monkeypatch.setattr(ops, "tuple_attention",
                            lambda *a, **k: (calls.append(1) if k.get("gather") is not None else None, orig(*a, **k))[1])
        g = batch([S.trimer_molecule(seed=s, n_confs=3) for s in range(3)]).to(DEV)
        loss, g = train_loss(model, g)
        loss.backward()
        monkeypatch.setattr(ops, "tuple_attention", orig)
        assert (len(calls) > 0) == dedup                      # bonds, angles and propers qualify at this size
        results[dedup] = (graph_params(g), float(loss), {n: p.grad.clone() for n, p in model.named_parameters() if p.grad is not None})
    for lvl in LEVELS:
        for key, v in results[False][0][lvl].items():
            assert torch.equal(v, results[True][0][lvl][key]), f"{lvl} {key}"
    assert results[False][1] == results[True][1]
    for n, gr in results[False][2].items():
        scale = float(gr.abs().max()) + 1e-12
        err = float((gr - results[True][2][n]).abs().max()) / scale
        assert err < 2e-4, f"{n}: {err:.2e}"


def test_drop_in_on_a_dgl_style_graph_object(golden_dir):
    """The modules only touch the accessor surface SURVEY.md §8b lists (ntypes, num_nodes, nodes[t].data, edges(etype),
    batch_num_nodes, to), so they must work on a DGL heterograph as well as on this package's container.  Real DGL is not
    installable here; the test infrastructure's pure-PyTorch stand-in with DGL's API (oracle/dgl_standin, the same object
    the unmodified reference ran on when the goldens were made) takes its place, for one molecule and for `dgl.batch`."""
    import sys
    standin = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle", "dgl_standin")
    if standin not in sys.path:
        sys.path.insert(0, standin)
    import dgl
    assert "standin" in dgl.__version__
    from make_golden import to_reference_graph
    gold = np.load(os.path.join(golden_dir, "cfg1_forward.npz"))
    model, _ = build_model()
    en = Energy(terms=["bond", "angle", "proper", "improper"])
    g = to_reference_graph(S.cfg1_graph(n_confs=5, seed=0)).to(DEV)
    assert type(g).__module__.startswith("dgl")
    with torch.no_grad():
        out = en(model(g))
    assert out is g                                           # same object, mutated (models/grappa.py:189)
    want = {l: {k: gold[f"{l}_{k}"] for k in (("k", "eq") if l in ("n2", "n3") else ("k",))} for l in LEVELS}
    assert_params_close(graph_params(g), want, "golden, dgl-style graph")
    np.testing.assert_allclose(g.nodes["g"].data["energy"].cpu().numpy(), gold["g_energy"], rtol=1e-4, atol=5e-2)
    np.testing.assert_allclose(g.nodes["n1"].data["gradient"].cpu().numpy(), gold["gradient"], rtol=1e-4, atol=2e-2)
    # a batch made by the DGL-style `batch` (idxs offset by the caller as utils/dgl_utils.py:57-59 does)
    mols = [S.trimer_molecule(seed=s, n_confs=3) for s in (0, 1)]
    ours = batch(mols).to(DEV)
    refs = [to_reference_graph(m) for m in mols]
    off = 0
    for r in refs:
        for lvl in LEVELS:
            r.nodes[lvl].data["idxs"] = r.nodes[lvl].data["idxs"] + off
        off += r.num_nodes("n1")
    gb = dgl.batch(refs).to(DEV)
    with torch.no_grad():
        a, b = en(model(ours)), en(model(gb))
    for lvl in LEVELS:
        for key, v in graph_params(a)[lvl].items():
            assert torch.equal(v, graph_params(b)[lvl][key]), f"{lvl} {key}"
    assert torch.equal(a.nodes["g"].data["energy"], b.nodes["g"].data["energy"])
    assert torch.equal(a.nodes["n1"].data["gradient"], b.nodes["n1"].data["gradient"])


def test_input_prefetch_and_async_loss_read_match_the_plain_step():
    """TrainStep.prefetch (H2D of the next batch on a copy stream into staging buffers) + read_loss_async give the same
    losses and weights as loading synchronously, with two different batches alternating."""
    from grappa_b200.batching import PackedBatch
    from grappa_b200.train_step import TrainStep
    batches = [PackedBatch(batch([S.trimer_molecule(seed=s + 10 * b, n_confs=4) for s in range(3)])) for b in range(2)]
    outs = []
    for pipelined in (False, True):
        model, _ = build_model(train=True)
        ops.DropoutSeeds.manual_seed(99)
        ts = TrainStep(model, DEV, use_cuda_graph=True)
        losses = []
        if pipelined:
            ts.prefetch(batches[0])
            pending = None
            for k in range(6):
                loss = ts.step(batches[k % 2])
                if k + 1 < 6:
                    ts.prefetch(batches[(k + 1) % 2])
                h = ts.read_loss_async(loss)
                if pending is not None:
                    losses.append(pending.value())
                pending = h
            losses.append(pending.value())
        else:
            for k in range(6):
                losses.append(float(ts.step(batches[k % 2])))
        torch.cuda.synchronize()
        outs.append((losses, ts.flat.params.clone()))
    assert outs[0][0] == outs[1][0]
    assert torch.equal(outs[0][1], outs[1][1])


def test_split_update_under_the_next_gnn_forward_is_bit_identical():
    """TrainStep(split_update=True) updates the writers' parameters on an update stream while the next step's GNN forward
    runs (an external-event wait node in the captured graph guards the writers): same losses, same weights."""
    from grappa_b200.batching import PackedBatch
    from grappa_b200.train_step import TrainStep
    packed = PackedBatch(batch([S.trimer_molecule(seed=s, n_confs=4) for s in range(3)]))
    outs = []
    for split in (False, True):
        model, _ = build_model(train=True)
        ops.DropoutSeeds.manual_seed(7)
        ts = TrainStep(model, DEV, use_cuda_graph=True, split_update=split)
        losses = [float(ts.step(packed)) for _ in range(7)]
        ts.synchronize()
        outs.append((losses, ts.flat.params.clone(), ts.flat.exp_avg_sq.clone()))
    assert outs[0][0] == outs[1][0]
    assert torch.equal(outs[0][1], outs[1][1]) and torch.equal(outs[0][2], outs[1][2])


def test_shape_cache_is_bounded_and_evicted_shapes_still_train():
    """One static buffer set + captured graph per batch shape, least recently used evicted beyond `max_cached_shapes`;
    a shape that comes back after eviction is simply warmed up and captured again."""
    from grappa_b200.batching import PackedBatch
    from grappa_b200.train_step import TrainStep
    shapes = [PackedBatch(batch([S.trimer_molecule(seed=s, n_confs=c) for s in range(n)])) for n, c in ((2, 3), (3, 3), (2, 5))]
    model, _ = build_model(train=True)
    ts = TrainStep(model, DEV, use_cuda_graph=True, max_cached_shapes=2)
    for rnd in range(2):
        for pk in shapes:
            for _ in range(5):                      # 3 warm-ups, capture, replay
                loss = ts.step(pk)
            assert torch.isfinite(loss)
            assert len(ts._static) <= 2
    assert shapes[0].signature not in ts._static and shapes[2].signature in ts._static


def test_folded_symmetriser_matches_the_explicit_permutations(monkeypatch):
    """Inference folds the symmetriser's permutations and the first LayerNorm's affine into weight views (shared
    statistics per tuple): same parameters as the explicit permuted copies, for both model configurations, and the fold
    follows weight updates."""
    import grappa_b200.model as M
    g_host = batch([S.trimer_molecule(seed=s, n_confs=2) for s in range(3)] +
                   [S.make_molecule(S.template_ace_ala_nme(), seed=4, n_confs=2, with_reference_data=True)])
    for make in (lambda: build_model()[0], lambda: GrappaModel().to(DEV).eval()):
        model = make()
        outs = []
        for fold in (False, True, True):
            monkeypatch.setattr(M, "FOLD_SYMMETRISER", fold)
            if len(outs) == 2:                        # change the weights in place: the cached fold must be rebuilt
                with torch.no_grad():
                    for p in model.parameters():
                        p.mul_(1.01)
                monkeypatch.setattr(M, "FOLD_SYMMETRISER", False)
                with torch.no_grad():
                    ref2 = graph_params(model(g_host.to(DEV)))
                monkeypatch.setattr(M, "FOLD_SYMMETRISER", True)
            with torch.no_grad():
                outs.append(graph_params(model(g_host.to(DEV))))
        # two evaluation orders of the same function: the bar is the rounding noise of the GEMM math mode in use
        # (3xTF32: 2e-5; bf16x3: 16 mantissa bits per operand, 4x that), far inside the 1e-4 parity budget either way
        tol = 2e-5 if ops.MATH_MODE < 2 else 8e-5
        for lvl in LEVELS:
            for key, v in outs[0][lvl].items():
                scale = float(v.abs().max()) + 1e-12
                assert float((v - outs[1][lvl][key]).abs().max()) / scale < tol, f"{lvl} {key}"
                assert float((ref2[lvl][key] - outs[2][lvl][key]).abs().max()) / (float(ref2[lvl][key].abs().max()) + 1e-12) < tol
