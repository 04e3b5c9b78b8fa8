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


def decode_energy_program(p, g):
    """Interprets the energy plan (csrc/energy.cu: block rows + per-block programs) back into tuples per block."""
    es = p.energy_sizes
    W = es["n_parts"]
    blk = p.host_view("e_blk").reshape(-1, 12)
    prog = p.host_view("e_prog")
    toff = p.host_view("e_mol_toff").reshape(4, -1)
    out = []
    for b in range(len(blk)):
        a0, na, mol, p0, nw, n_slots, nb, j, local = blk[b, :9]
        assert p0 % 4 == 0 and nw % 4 == 0
        pg = prog[p0:p0 + nw]
        offs = pg[:W + 1]
        pull_ptr = pg[W + 1:W + 2 + na]
        pull = pg[W + 2 + na:W + 2 + na + pull_ptr[-1]]
        assert pull_ptr[0] == 0 and (np.diff(pull_ptr) >= 0).all()
        smap = np.full((W, na), -1, np.int64)                       # [part][atom] -> 96 * slot, from the pull lists
        for al in range(na):
            ent = pull[pull_ptr[al]:pull_ptr[al + 1]]
            assert (np.diff(ent >> 16) > 0).all()                    # part order = summation order
            smap[ent >> 16, al] = ent & 0xffff
        assert offs[0] == W + 2 + na + pull_ptr[-1] and offs[-1] <= nw
        recs = []       # (part, level, tuple row, owner, atom rows)
        for w in range(W):
            pos = offs[w]
            counts = pg[pos:pos + 4]
            pos += 4
            used = smap[w][smap[w] >= 0]
            assert (used % 96 == 0).all() and len(set(used.tolist())) == len(used) and (used // 96 < n_slots).all()
            touched = set()
            for level, n, ns in ((2, counts[0], 4), (3, counts[1], 4), (1, counts[2], 3), (0, counts[3], 2)):
                for _ in range(n):
                    hdr = int(pg[pos])
                    words = pg[pos + 1:pos + 1 + ns].astype(np.int64)
                    if local:
                        al, slot = (words & 0xffff), (words >> 16) & 0xffff
                        assert (al % 384 == 0).all() and (slot % 384 == 0).all()
                        al, slot = al // 384, slot // 384
                        assert (smap[w][al] == slot * 96).all()      # the reduction map names the same slots
                        atoms = a0 + al
                    else:
                        al = words
                        atoms = a0 + words
                        assert (smap[w] == 96 * np.arange(na)).all()
                    touched.update(int(x) for x in al if 0 <= x < na)
                    recs.append((w, level, int(toff[level, mol]) + (hdr >> 1), hdr & 1, atoms.tolist()))
                    pos += 1 + ns
            if local:
                assert touched == set(np.nonzero(smap[w] >= 0)[0].tolist())
            assert pos == offs[w + 1]
        out.append(recs)
    return blk, out


@pytest.mark.parametrize("block_atoms,n_parts", [(64, 4), (16, 3), (7, 1), (64, 8), (64, 16)])
def test_energy_plan_invariants(block_atoms, n_parts):
    g = batch([S.trimer_molecule(s, 2) for s in (0, 1, 2)])
    p = build_plan(g, block_atoms=block_atoms, n_parts=n_parts)
    es = p.energy_sizes
    blk, recs = decode_energy_program(p, g)
    n = g.num_nodes("n1")
    # blocks partition the atoms, never cross molecules, and respect the size limit
    assert blk[:, 1].sum() == n and (blk[:, 1] <= block_atoms).all()
    atom_blk = np.repeat(np.arange(len(blk)), blk[:, 1])
    mol_of_atom = np.repeat(np.arange(3), 40)
    assert (mol_of_atom == blk[atom_blk, 2]).all()
    assert es["any_multi"] == es["any_nonlocal"] == int(block_atoms < 40)
    assert es["n_parts"] == n_parts and es["n_blocks"] == len(blk) and es["n_words"] == len(p.host_view("e_prog"))
    assert es["max_words"] == blk[:, 4].max() and es["max_atoms"] == blk[:, 1].max()
    levels = ("n2", "n3", "n4", "n4_improper")
    assert [es["max_t2"], es["max_t3"], es["max_t4"], es["max_t4i"]] == [int(g.batch_num_nodes(l).max()) for l in levels]
    owners = {l: np.zeros(g.num_nodes(l), int) for l in levels}
    seen = {l: [set() for _ in range(len(blk))] for l in levels}
    for b in range(len(blk)):
        per_kind = {}
        for w, level, t, owner, atoms in recs[b]:
            idx = g.nodes[levels[level]].data["idxs"].numpy()[t]
            assert atoms == idx.tolist()                       # the record names the tuple's atoms, in order
            assert (atom_blk[idx] == b).any()                  # and the tuple touches this block
            assert blk[b, 8] == 0 or (atom_blk[idx] == b).all()
            assert t not in seen[levels[level]][b]
            seen[levels[level]][b].add(t)
            if owner:
                owners[levels[level]][t] += 1
                assert atom_blk[idx[0]] == b
            per_kind[w] = per_kind.get(w, 0) + {0: 6, 1: 14, 2: 18, 3: 18}[level]
        c = [per_kind.get(w, 0) for w in range(n_parts)]        # parts are balanced by estimated cost
        assert max(c) - min(c) <= 36
    for li, l in enumerate(levels):
        assert (owners[l] == 1).all()       # every tuple's energy is counted exactly once
        idx = g.nodes[l].data["idxs"].numpy()
        for t in range(len(idx)):           # and it is evaluated in exactly the blocks it touches
            assert {b for b in range(len(blk)) if t in seen[l][b]} == set(atom_blk[idx[t]].tolist())
    tm = p.host_view("tuple_mol")
    off = 0
    for l in levels:
        idx = g.nodes[l].data["idxs"].numpy()
        np.testing.assert_array_equal(tm[off:off + len(idx)], mol_of_atom[idx[:, 0]])
        off += len(idx)
    toff = p.host_view("e_mol_toff").reshape(4, -1)
    for li, l in enumerate(levels):
        np.testing.assert_array_equal(toff[li], np.concatenate([[0], np.cumsum(g.batch_num_nodes(l).numpy())]))


def test_energy_plan_rejects_ungrouped_tuples():
    g = batch([S.trimer_molecule(s, 2) for s in (0, 1)])
    ix = g.nodes["n2"].data["idxs"].clone()
    g.nodes["n2"].data["idxs"] = torch.flip(ix, dims=[0])      # molecule 1's bonds first
    with pytest.raises(RuntimeError, match="grouped by molecule"):
        build_plan(g)


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
    assert p.T["n4_improper"] == 0 and p.energy_sizes["max_t4i"] == 0
    _, recs = decode_energy_program(p, g)
    assert len(recs) == 1 and len(recs[0]) == 21 + 36 + 41
