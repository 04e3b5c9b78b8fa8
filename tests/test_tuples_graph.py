"""Tuple enumeration and batching layouts are bit-exact against golden vectors produced by the reference's
utils/tuple_indices.py and utils/dgl_utils.py (and against the live reference when its tree is present)."""
import os

import numpy as np
import pytest
import torch

from grappa_b200 import synthetic as S
from grappa_b200.graph import batch, unbatch
from grappa_b200.tuples import enumerate_tuples, enumerate_tuples_numpy, improper_rows


def test_tuples_match_golden(golden_dir):
    gold = np.load(os.path.join(golden_dir, "tuples.npz"))
    for name in ("ace", "AFA", "poly20", "shuffledFAF"):
        bonds = gold[name + "_bonds"]
        n = int(bonds.max()) + 1
        t = enumerate_tuples(bonds, n)
        np.testing.assert_array_equal(t["angles"], gold[name + "_angles"])
        np.testing.assert_array_equal(t["propers"], gold[name + "_propers"])
        nb = S._neighbour_lists(bonds, n)
        cen = gold[name + "_centres"]
        np.testing.assert_array_equal(improper_rows(cen, [nb[c] for c in cen.tolist()]), gold[name + "_impropers"])


def test_cxx_and_numpy_enumerations_agree_on_random_graphs():
    rng = np.random.default_rng(0)
    for trial in range(20):
        n = int(rng.integers(3, 60))
        # random tree plus a few ring-closing bonds, atoms relabelled so that first-appearance order != index order
        perm = rng.permutation(n)
        bonds = [(perm[i], perm[int(rng.integers(0, i))]) for i in range(1, n)]
        have = {tuple(sorted(b)) for b in bonds}
        for _ in range(int(rng.integers(0, 4))):
            a, b = (int(x) for x in rng.choice(n, 2, replace=False))
            if (min(a, b), max(a, b)) not in have:
                have.add((min(a, b), max(a, b)))
                bonds.append((a, b))
        bonds = np.array(bonds, np.int64)[rng.permutation(len(bonds))]
        a, b = enumerate_tuples(bonds, n), enumerate_tuples_numpy(bonds, n)
        for k in ("bonds", "angles", "propers"):
            np.testing.assert_array_equal(a[k], b[k], err_msg=f"trial {trial} {k}")
        # every angle / proper is a path along bonds, each equivalence class appears once
        assert len({tuple(r) for r in a["angles"].tolist()} | {tuple(r[::-1]) for r in a["angles"].tolist()}) == 2 * len(a["angles"])
        for r in a["propers"].tolist():
            assert all((min(x, y), max(x, y)) in have for x, y in zip(r[:-1], r[1:])) and r[0] < r[3]


def test_survey_sizes():
    g = S.cfg1_graph(n_confs=2)
    assert [g.num_nodes(t) for t in ("n1", "n2", "n3", "n4", "n4_improper")] == [22, 21, 36, 41, 12]
    g = S.trimer_molecule(0, 2)
    assert [g.num_nodes(t) for t in ("n1", "n2", "n3", "n4", "n4_improper")] == [40, 40, 68, 95, 24]
    g = S.cfg3_graph()
    assert [g.num_nodes(t) for t in ("n1", "n2", "n3", "n4", "n4_improper")] == [15000, 15499, 26996, 39990, 11997]
    assert g.num_edges() == 30998


def test_empty_and_edge_cases():
    t = enumerate_tuples(np.zeros((0, 2), np.int64), 0)
    assert t["angles"].shape == (0, 3) and t["propers"].shape == (0, 4)
    t = enumerate_tuples([[0, 1]], 2)           # a single bond: no angles, no propers
    assert len(t["angles"]) == 0 and len(t["propers"]) == 0
    t = enumerate_tuples([[0, 1], [1, 2], [2, 0]], 3)    # 3-ring: 3 angles, no propers
    assert len(t["angles"]) == 3 and len(t["propers"]) == 0
    assert improper_rows([], []).shape == (0, 4)
    with pytest.raises(ValueError):
        enumerate_tuples([[1, 1]], 2)


def test_batch_layout_matches_reference_golden(golden_dir):
    gold = np.load(os.path.join(golden_dir, "batch3_train.npz"))
    mols = [S.trimer_molecule(seed=s, n_confs=4) for s in (0, 1, 2)]
    before = [m.nodes["n4"].data["idxs"].clone() for m in mols]
    g = batch(mols)
    for lvl in ("n2", "n3", "n4", "n4_improper"):
        np.testing.assert_array_equal(g.nodes[lvl].data["idxs"].numpy(), gold[f"batched_{lvl}_idxs"])
        np.testing.assert_array_equal(g.batch_num_nodes(lvl).numpy(), gold[f"batch_num_{lvl}"])
    np.testing.assert_array_equal(torch.stack(g.edges()).numpy(), gold["batched_edges"])
    for m, b in zip(mols, before):           # inputs are not mutated (deliberate difference to dgl_utils.py:57-59)
        assert torch.equal(m.nodes["n4"].data["idxs"], b)


def test_unbatch_roundtrip_and_dummy_confs():
    mols = [S.trimer_molecule(seed=s, n_confs=4) for s in (3, 4)]
    mols[1].nodes["g"].data["is_dummy"] = torch.tensor([[0., 0., 1., 1.]])
    g = batch(mols)
    g.nodes["g"].data["energy"] = torch.arange(8.).view(2, 4)
    g.nodes["n1"].data["gradient"] = torch.randn(80, 4, 3)
    parts = unbatch(g)
    assert len(parts) == 2
    for p, m in zip(parts, mols):
        for lvl in ("n2", "n3", "n4", "n4_improper"):
            assert torch.equal(p.nodes[lvl].data["idxs"], m.nodes[lvl].data["idxs"])
        assert torch.equal(torch.stack(p.edges()), torch.stack(m.edges()))
    assert parts[0].nodes["n1"].data["xyz"].shape[1] == 4
    assert parts[1].nodes["n1"].data["xyz"].shape[1] == 2
    assert parts[1].nodes["g"].data["energy"].tolist() == [[4., 5.]]
    assert parts[1].nodes["n1"].data["gradient"].shape == (40, 2, 3)


def test_batch_rejects_mixed_conformation_counts():
    with pytest.raises(ValueError):
        batch([S.trimer_molecule(0, 4), S.trimer_molecule(1, 5)])


@pytest.mark.reference
def test_tuples_match_live_reference():
    import reference_loader as rl
    if not rl.reference_available():
        pytest.skip("reference tree not present")
    rl.load_reference()
    from grappa.utils import tuple_indices as TI
    rng = np.random.default_rng(7)
    for seq in ("AF", "FFA", "AFAFAF"):
        t = S.template_polymer(seq)
        bonds = t["bonds"][rng.permutation(len(t["bonds"]))]
        flip = rng.random(len(bonds)) < 0.5
        bonds[flip] = bonds[flip][:, ::-1]
        bl = [tuple(b) for b in bonds.tolist()]
        nd = TI.get_neighbor_dict(bl, sort=True)
        ref = TI.get_idx_tuples(bl, neighbor_dict=nd, is_sorted=True)
        mine = enumerate_tuples(bonds, len(t["atomic_numbers"]))
        np.testing.assert_array_equal(mine["angles"], np.array(ref["angles"]).reshape(-1, 3))
        np.testing.assert_array_equal(mine["propers"], np.array(ref["propers"]).reshape(-1, 4))
