"""Real molecules from the reference tree's only binary fixture (data_old/dgl_datasets/small_example/graphs.bin: 12 small
molecules with 40-50 conformations, QM energies / forces), read with this repo's DGL-free reader
and committed as tests/golden/small_example.npz.  They add what the synthetic peptides lack: varied elements and
degrees, rings, molecules WITHOUT proper torsions or impropers (empty levels), ragged conformation counts."""
import os
import struct

import numpy as np
import pytest
import torch

import grappa_oracle as O
from weights import deterministic_fill_
from real_data import LEVELS, load_real_molecules
from grappa_b200 import dgl_bin
from grappa_b200.graph import batch, set_number_confs
from grappa_b200.tuples import enumerate_tuples

REF_FILE = "/root/reference/data_old/dgl_datasets/small_example/graphs.bin"


def test_tuple_rows_of_real_molecules_equal_the_enumeration(golden_dir):
    """The angle / proper rows stored in the file were produced by the reference pipeline (utils/tuple_indices.py via
    data/molecule.py): an independent pin of grappa_enumerate_tuples' row ORDER on real topologies."""
    z = np.load(os.path.join(golden_dir, "small_example.npz"))
    n_with_propers = 0
    for i in range(int(z["n_graphs"])):
        bonds = z[f"g{i}_n2_idxs"]
        n = z[f"g{i}_n1_xyz"].shape[0]
        t = enumerate_tuples(bonds, n)
        np.testing.assert_array_equal(t["angles"], z[f"g{i}_n3_idxs"], err_msg=f"graph {i} angles")
        np.testing.assert_array_equal(t["propers"], z[f"g{i}_n4_idxs"], err_msg=f"graph {i} propers")
        n_with_propers += len(t["propers"]) > 0
        imp = z[f"g{i}_n4_improper_idxs"]
        if len(imp):                                   # three versions per centre, central atom at position 2, all bonded to it
            assert len(imp) % 3 == 0
            bonded = {tuple(sorted(b)) for b in bonds.tolist()}
            for row in imp.tolist():
                assert all(tuple(sorted((row[2], row[j]))) in bonded for j in (0, 1, 3))
    assert n_with_propers >= 6


def _write_dgl_file(path, graphs):
    """Minimal writer of the layout dgl_bin.load_graphs documents (structure section replaced by filler bytes)."""
    def s(x):
        b = x.encode()
        return struct.pack("<Q", len(b)) + b

    def nd(a):
        a = np.ascontiguousarray(a)
        code = {"i": 0, "u": 1, "f": 2}[a.dtype.kind]
        return (struct.pack("<QQiii", dgl_bin.NDARRAY_MAGIC, 0, 1, 0, a.ndim) + struct.pack("<BBH", code, a.dtype.itemsize * 8, 1)
                + struct.pack(f"<{a.ndim}q", *a.shape) + struct.pack("<q", a.nbytes) + a.tobytes())

    out = struct.pack("<4Q", dgl_bin.FILE_MAGIC, 2, 2, len(graphs)).ljust(4096, b"\xab")
    out += struct.pack("<2Q", 4112, 0)
    for g in graphs:
        out += struct.pack("<Q", dgl_bin.HETERO_MAGIC) + bytes(range(200))          # "graph structure"
        ntypes = sorted(g)
        out += struct.pack("<Q", len(ntypes))
        for nt in ntypes:
            out += struct.pack("<Q", len(g[nt]))
            for k, v in g[nt].items():
                out += s(k) + nd(v)
        out += struct.pack("<Q", len(ntypes)) + b"".join(struct.pack("<Q", 0) for _ in ntypes)   # no edge tensors
        out += struct.pack("<Q", len(ntypes)) + b"".join(s(nt) for nt in ntypes)
        out += struct.pack("<Q", len(ntypes)) + b"".join(s(nt + "_edge") for nt in ntypes)
    with open(path, "wb") as f:
        f.write(out + b"trailer")


def test_dgl_bin_reader_round_trip(tmp_path):
    rng = np.random.default_rng(0)
    graphs = []
    for n in (3, 7):
        graphs.append({"n1": {"xyz": rng.normal(size=(n, 4, 3)).astype(np.float32), "ids": np.arange(n, dtype=np.int64),
                              "atomic_number": np.eye(53, dtype=np.float32)[rng.integers(0, 53, n)]},
                       "n2": {"idxs": np.stack([np.arange(n - 1), np.arange(1, n)], 1).astype(np.int64)},
                       "n3": {"idxs": np.zeros((0, 3), np.int64)}, "g": {"energy_qm": rng.normal(size=(1, 4)).astype(np.float32)}})
    path = str(tmp_path / "graphs.bin")
    _write_dgl_file(path, graphs)
    got = dgl_bin.load_graphs(path)
    assert len(got) == 2
    for a, b in zip(got, graphs):
        assert sorted(a) == sorted(b)
        for nt in b:
            for k, v in b[nt].items():
                assert a[nt][k].dtype == v.dtype and np.array_equal(a[nt][k], v), (nt, k)
    hg = dgl_bin.to_heterograph(got[1])
    assert hg.num_nodes("n1") == 7 and hg.num_edges() == 12 and hg.num_nodes("n3") == 0
    with open(path, "r+b") as f:                      # corrupt the magic
        f.write(b"\x00")
    with pytest.raises(ValueError):
        dgl_bin.load_graphs(path)


@pytest.mark.reference
def test_reader_on_the_reference_fixture_matches_the_committed_npz(golden_dir):
    if not os.path.exists(REF_FILE):
        pytest.skip("reference tree not present")
    graphs = dgl_bin.load_graphs(REF_FILE)
    z = np.load(os.path.join(golden_dir, "small_example.npz"))
    assert len(graphs) == int(z["n_graphs"]) == 12
    for i, g in enumerate(graphs):
        assert sorted(g) == ["g", "n1", "n2", "n3", "n4", "n4_improper"]
        np.testing.assert_array_equal(g["n1"]["xyz"], z[f"g{i}_n1_xyz"])
        np.testing.assert_array_equal(g["g"]["energy_qm"], z[f"g{i}_g_energy_qm"])
        np.testing.assert_array_equal(np.asarray(g["n4"]["idxs"]).reshape(-1, 4), z[f"g{i}_n4_idxs"])
        n = g["n1"]["xyz"].shape[0]
        assert g["n1"]["atomic_number"].shape == (n, 53) and np.all(g["n1"]["atomic_number"].sum(1) == 1)
        assert int(np.asarray(g["n2"]["idxs"]).max()) == n - 1 or n == 1


@pytest.mark.gpu
def test_real_molecules_forward_energy_and_training_gradients_match_oracle(golden_dir):
    from grappa_b200 import ops
    from grappa_b200.energy import Energy
    from grappa_b200.model import GrappaModel
    from grappa_b200.plan import get_plan
    mols = [set_number_confs(m, 8) for m in load_real_molecules(golden_dir)]        # ragged 40-50 -> 8 (random subset)
    g = batch(mols)
    model = GrappaModel.from_default_config()
    sd = model.state_dict()
    deterministic_fill_(sd)
    model.load_state_dict(sd)
    model = model.to("cuda").train()
    for m in model.modules():
        if isinstance(m, torch.nn.Dropout):
            m.p = 0.0
    fixed = ("to_k.", "to_eq.", "k_mean", "k_std", "prefactors", "positional")
    leaf = {k: (v.clone().requires_grad_(True) if torch.is_floating_point(v) and not any(s in k for s in fixed) else v.clone())
            for k, v in sd.items()}
    assert torch.isnan(O.train_step_loss(sd, g)[0])          # the reference's loss is NaN here (molecules without impropers)
    loss_ref, g_ref, _ = O.train_step_loss(leaf, g, nan_free=True)
    loss_ref.backward()
    gd = torch.nn.Sequential(model, Energy(terms=["bond", "angle", "proper", "improper"]))(g.to("cuda"))
    plan = get_plan(gd)
    loss = ops.MolwiseLossFn.apply(gd.nodes["g"].data["energy"], gd.nodes["n1"].data["gradient"], gd.nodes["n4"].data["k"],
                                   gd.nodes["n4_improper"].data["k"], gd.nodes["g"].data["energy_ref"],
                                   gd.nodes["n1"].data["gradient_ref"], gd.nodes["g"].data.get("is_dummy"), plan, (1.0, 0.8, 1.0, 1.0))
    loss.backward()
    assert gd.nodes["n4"].data["k"].shape == (g.num_nodes("n4"), 3)
    np.testing.assert_allclose(loss.item(), loss_ref.item(), rtol=2e-4)
    worst = 0.0
    for name, p in model.named_parameters():
        r = leaf[name].grad
        if r is None:
            assert p.grad is None, name
            continue
        scale = float(r.abs().max())
        np.testing.assert_allclose(p.grad.cpu().numpy(), r.numpy(), rtol=2e-3, atol=3e-4 * scale, err_msg=name)
        worst = max(worst, float((p.grad.cpu() - r).abs().max()) / (scale + 1e-30))
    print("worst relative gradient error on real molecules:", worst)
