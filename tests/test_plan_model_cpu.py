"""Host-side logic: batch plan structure (CSR, incidence, energy blocks) and the state-dict drop-in contract."""
import json
import os

import numpy as np
import pytest
import torch

from grappa_b200 import synthetic as S
from grappa_b200.graph import batch
from grappa_b200.model import GrappaModel
from grappa_b200.plan import build_plan


def test_state_dict_layout_matches_reference(golden_dir):
    with open(os.path.join(golden_dir, "state_dict_layout.json")) as f:
        layout = json.load(f)
    sd = GrappaModel.from_default_config().state_dict()
    assert list(sd.keys()) == list(layout.keys())
    assert len(sd) == 296
    for k, (shape, dtype) in layout.items():
        assert list(sd[k].shape) == shape, k
        assert str(sd[k].dtype) == dtype, k
    n_params = sum(p.numel() for p in GrappaModel.from_default_config().parameters())
    assert n_params == 20836867


def test_model_attributes_and_unsupported_variants():
    m = GrappaModel.from_default_config()
    assert m.field_of_view == 7
    assert m.gnn.blocks[0] is m.gnn.att_blocks[0]
    seq = torch.nn.Sequential(m, torch.nn.Identity())
    assert seq[0] is m
    with pytest.raises(NotImplementedError):
        GrappaModel.from_default_config(gnn_convolutions=1)
    with pytest.raises(NotImplementedError):
        GrappaModel.from_default_config(harmonic_gate=True)
    py_default = GrappaModel()          # the Python defaults differ from the YAML (models/grappa.py:28-29)
    assert py_default.parameter_writer.proper_writer.gated is False
    assert int(py_default.parameter_writer.improper_writer.n_periodicity) == 3


def _check_csr(ptr, nbr, keys, other, n):
    assert ptr[0] == 0 and ptr[-1] == len(keys)
    for v in range(n):
        mine = nbr[ptr[v]:ptr[v + 1]].tolist()
        ref = other[keys == v].tolist()     # stable: original edge order
        assert mine == ref


def test_plan_csr_and_incidence():
    g = batch([S.trimer_molecule(s, 3) for s in (0, 1)])
    p = build_plan(g)
    n = g.num_nodes("n1")
    src, dst = (t.numpy().astype(np.int64) for t in g.edges())
    _check_csr(p.host_view("in_ptr"), p.host_view("in_nbr"), dst, src, n)
    _check_csr(p.host_view("out_ptr"), p.host_view("out_nbr"), src, dst, n)
    for lvl, s in (("n2", 2), ("n3", 3), ("n4", 4), ("n4_improper", 4)):
        idx = g.nodes[lvl].data["idxs"].numpy()
        np.testing.assert_array_equal(p.host_view("idx_" + lvl).reshape(-1, s), idx)
        ptr, tok = p.host_view("inc_ptr_" + lvl), p.host_view("inc_tok_" + lvl)
        flat = idx.reshape(-1)
        for a in range(n):
            assert tok[ptr[a]:ptr[a + 1]].tolist() == np.nonzero(flat == a)[0].tolist()
    np.testing.assert_array_equal(p.host_view("mol_atom_off"), [0, 40, 80])
    np.testing.assert_array_equal(p.host_view("mol_off_n4"), np.concatenate([[0], np.cumsum(g.batch_num_nodes("n4").numpy())]))


@pytest.mark.parametrize("block_atoms", [64, 16, 7])
def test_energy_plan_invariants(block_atoms):
    g = batch([S.trimer_molecule(s, 2) for s in (0, 1, 2)])
    p = build_plan(g, block_atoms=block_atoms)
    es = p.energy_sizes
    blk = p.host_view("e_blk").reshape(-1, 12)
    rec_atoms = p.host_view("e_rec_atoms").reshape(-1, 4)
    rec_meta = p.host_view("e_rec_meta").reshape(-1, 2)
    rec_slot = p.host_view("e_rec_slot").reshape(-1, 4)
    pull_ptr, pull_slot = p.host_view("e_pull_ptr"), p.host_view("e_pull_slot")
    n = g.num_nodes("n1")
    # blocks partition the atoms, never cross molecules, and respect the size limit
    assert blk[:, 1].sum() == n and (blk[:, 1] <= block_atoms).all()
    atom_blk = np.repeat(np.arange(len(blk)), blk[:, 1])
    mol_of_atom = np.repeat(np.arange(3), 40)
    assert (mol_of_atom == blk[atom_blk, 2]).all()
    assert es["any_multi"] == int(block_atoms < 40)
    n_seq = {0: 2, 1: 3, 2: 4, 3: 4}
    levels = ("n2", "n3", "n4", "n4_improper")
    owners = {l: np.zeros(g.num_nodes(l), int) for l in levels}
    total_inc = 0
    for b in range(len(blk)):
        r0, nr = blk[b, 3], blk[b, 4]
        seen_slots = []
        prev = (-1, -1)
        for r in range(r0, r0 + nr):
            level, owner, t = rec_meta[r, 0] & 3, (rec_meta[r, 0] >> 2) & 1, rec_meta[r, 1]
            assert (level, t) > prev        # ordered by (level, tuple) inside a block
            prev = (level, t)
            idx = g.nodes[levels[level]].data["idxs"].numpy()[t]
            np.testing.assert_array_equal(rec_atoms[r, :n_seq[level]], idx)
            if owner:
                owners[levels[level]][t] += 1
                assert atom_blk[idx[0]] == b
            for i in range(4):
                inside = i < n_seq[level] and atom_blk[idx[i]] == b
                assert (rec_slot[r, i] >= 0) == inside
                if inside:
                    seen_slots.append(rec_slot[r, i])
                    a = idx[i]
                    assert rec_slot[r, i] in pull_slot[pull_ptr[a]:pull_ptr[a + 1]]
        assert sorted(seen_slots) == list(range(blk[b, 5]))     # slots are dense per block
        lv = rec_meta[r0:r0 + nr, 0] & 3
        assert [int((lv == l).sum()) for l in range(4)] == blk[b, 8:12].tolist()
        total_inc += len(seen_slots)
    for l in levels:
        assert (owners[l] == 1).all()       # every tuple's energy is counted exactly once
    assert total_inc == es["n_pull"] == pull_ptr[-1] == sum(g.num_nodes(l) * s for l, s in zip(levels, (2, 3, 4, 4)))
    tm = p.host_view("tuple_mol")
    off = 0
    for l in levels:
        idx = g.nodes[l].data["idxs"].numpy()
        np.testing.assert_array_equal(tm[off:off + len(idx)], mol_of_atom[idx[:, 0]])
        off += len(idx)


def test_plan_rejects_out_of_range_idxs():
    g = S.cfg1_graph(n_confs=1)
    g.nodes["n3"].data["idxs"] = g.nodes["n3"].data["idxs"].clone()
    g.nodes["n3"].data["idxs"][0, 0] = 22
    with pytest.raises(AssertionError, match="Encountered idxs up to 22"):
        build_plan(g)


def test_plan_handles_missing_and_empty_levels():
    g = S.cfg1_graph(n_confs=1)
    g.nodes["n4_improper"].data["idxs"] = torch.zeros((0, 4), dtype=torch.long)
    g._num_nodes["n4_improper"] = 0
    g._batch_num_nodes["n4_improper"] = torch.tensor([0])
    p = build_plan(g)
    assert p.T["n4_improper"] == 0 and p.energy_sizes["n_records"] == 21 + 36 + 41
