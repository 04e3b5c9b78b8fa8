"""Device-side collate: the segment tables that `MoleculeStore.prepare` emits, applied to the
packed per-molecule store, must reproduce bit for bit the host path
`PackedBatch(batch([set_number_confs(m, C, conf_idxs=...)]))`, which itself follows utils/dgl_utils.py:11-63,137-176
and data/graph_data_loader.py:23-73.  The CPU tests interpret the tables with a numpy restatement of the two gather
kernels; the GPU test runs the kernels themselves through the C ABI."""
import numpy as np
import pytest
import torch

from grappa_b200 import synthetic as S
from grappa_b200.batching import DeviceBatch, PackedBatch
from grappa_b200.graph import batch, conf_indices, set_number_confs
from grappa_b200.store import MoleculeStore


def apply_int_segments(seg, store, lit, out):
    for src, dst, ln, mode, a0, a1, a2, a3 in seg.tolist():
        i = np.arange(ln)
        if mode == 0:
            v = store[src:src + ln] + a0
        elif mode == 1:
            c = i % 12
            v = store[src:src + ln] + np.where(c == 0, a0, np.where(c == 2, a1, np.where(c == 3, a2, 0)))
        elif mode == 4:
            v = np.full(ln, a0)
        else:
            v = lit[src:src + ln]
        out[dst:dst + ln] = v


def apply_float_segments(seg, conf, store, out):
    for src, dst, rows, inner, cs, cd, coff, mode in seg.tolist():
        if mode == 1:
            out[dst:dst + cd] = (np.arange(cd) >= cs).astype(np.float32)
            continue
        block = store[src:src + rows * cs * inner].reshape(rows, cs, inner)
        sel = np.arange(cd) if coff < 0 else conf[coff:coff + cd]
        out[dst:dst + rows * cd * inner] = block[:, sel, :].reshape(-1)


def host_reference(mols, ids, conf_idxs, n_parts):
    from grappa_b200.plan import build_plan
    C = conf_idxs.shape[1]
    g = batch([set_number_confs(mols[i], C, conf_idxs=conf_idxs[j]) for j, i in enumerate(ids)])
    return PackedBatch(g, pin=False, plan=build_plan(g, n_parts=n_parts))      # same energy parts as the store's molecules


def make_molecules():
    """Different sizes and different numbers of stored conformations."""
    mols = [S.trimer_molecule(seed=s, n_confs=c) for s, c in ((0, 6), (1, 4), (2, 9), (3, 6))]
    mols.append(S.make_molecule(S.template_ace_ala_nme(), seed=7, n_confs=5, with_reference_data=True))
    # > 64 atoms: the molecule spans several energy blocks (records shared between blocks, cross-block slots)
    mols.append(S.make_molecule(S.template_polymer(S.alternating_sequence(8)), seed=1, n_confs=3, with_reference_data=True))
    return mols


def check_sections(spec, ibuf, fbuf, ref):
    assert spec.signature == ref.signature
    for name, (s, n) in ref.plan.offsets.items():
        np.testing.assert_array_equal(ibuf[s:s + n], ref.plan.host.numpy()[s:s + n], err_msg=name)
    for key, (off, shape) in ref.layout.items():
        n = int(np.prod(shape))
        np.testing.assert_array_equal(fbuf[off:off + n], ref.fbuf.numpy()[off:off + n], err_msg=str(key))


@pytest.mark.parametrize("ids,C", [([0, 1, 2], 4), ([2, 2, 0, 3], 6), ([4, 0], 7), ([1], 4), ([5, 0, 4], 5)])
def test_segment_tables_reproduce_host_collate(ids, C):
    mols = make_molecules()
    ids = [i for i in ids if i < len(mols)]
    store = MoleculeStore(mols, "cpu")
    gen = torch.Generator().manual_seed(3)
    conf = torch.stack([conf_indices(int(mols[i].nodes["n1"].data["xyz"].shape[1]), C, gen) for i in ids])
    spec = store.prepare(ids, conf_idxs=conf)
    ref = host_reference(mols, ids, conf, store.energy_parts)
    ibuf = np.full(spec.n_int, -77, np.int32)
    fbuf = np.full(spec.n_float, np.nan, np.float32)
    seg = spec.seg64.numpy()
    tab = spec.tab32.numpy()
    apply_int_segments(seg[:spec.n_int_seg], store.istore.numpy(), tab[spec.lit_off:], ibuf)
    apply_float_segments(seg[spec.n_int_seg:], tab, store.fstore.numpy(), fbuf)
    check_sections(spec, ibuf, fbuf, ref)
    # padded conformations are flagged, sub-sampled ones are not
    off, shape = spec.layout[("g", "is_dummy")]
    flags = fbuf[off:off + len(ids) * C].reshape(len(ids), C)
    for j, i in enumerate(ids):
        present = int(mols[i].nodes["n1"].data["xyz"].shape[1])
        np.testing.assert_array_equal(flags[j], (np.arange(C) >= present).astype(np.float32))


def test_conf_strategy_follows_reference_rule():
    mols = make_molecules()
    store = MoleculeStore(mols, "cpu")
    ids = np.array([0, 1, 2])                     # 6, 4, 9 stored conformations
    assert store.n_confs_for(ids, "min") == 4 and store.n_confs_for(ids, "max") == 9 and store.n_confs_for(ids, "all") == 9
    assert store.n_confs_for(ids, "mean") == 6 and store.n_confs_for(ids, 5) == 5 and store.n_confs_for(ids, 50) == 9
    with pytest.raises(ValueError):
        store.n_confs_for(ids, "median")
    with pytest.raises(IndexError):
        store.prepare([0, 99])
    with pytest.raises(ValueError):
        store.prepare([])
    spec = store.prepare(ids, "min", generator=torch.Generator().manual_seed(0))
    assert spec.conf_idxs.shape == (3, 4)
    assert sorted(spec.conf_idxs[1].tolist()) == [0, 1, 2, 3]          # nothing to drop: stored order kept
    assert spec.conf_idxs[1].tolist() == [0, 1, 2, 3]
    assert len(set(spec.conf_idxs[2].tolist())) == 4                    # random subset without repeats


def test_set_number_confs_matches_reference_semantics():
    g = S.trimer_molecule(seed=0, n_confs=3)
    xyz = g.nodes["n1"].data["xyz"].clone()
    out = set_number_confs(g, 5)
    assert out.nodes["g"].data["is_dummy"].tolist() == [[0, 0, 0, 1, 1]]
    assert torch.equal(out.nodes["n1"].data["xyz"][:, 3], xyz[:, 2]) and torch.equal(out.nodes["n1"].data["xyz"][:, 4], xyz[:, 2])
    assert torch.equal(g.nodes["n1"].data["xyz"], xyz)                   # input untouched
    out = set_number_confs(g, 2, generator=torch.Generator().manual_seed(1))
    assert out.nodes["n1"].data["xyz"].shape[1] == 2 and out.nodes["g"].data["energy_ref"].shape == (1, 2)
    assert out.nodes["g"].data["is_dummy"].sum() == 0


@pytest.mark.gpu
def test_device_collate_bit_exact_and_trainable():
    dev = torch.device("cuda:0")
    mols = make_molecules()
    store = MoleculeStore(mols, dev)
    for ids, C in (([0, 1, 2], 4), ([2, 2, 0, 3], 6), ([4, 0], 7), ([len(mols) - 1, 0, 4], 5)):
        gen = torch.Generator().manual_seed(5)
        conf = torch.stack([conf_indices(int(mols[i].nodes["n1"].data["xyz"].shape[1]), C, gen) for i in ids])
        ref = host_reference(mols, ids, conf, store.energy_parts)
        db = store.collate(ids, conf_idxs=conf)
        torch.cuda.synchronize()
        spec = store.prepare(ids, conf_idxs=conf)
        check_sections(spec, db.ibuf.cpu().numpy(), db.fbuf.cpu().numpy(), ref)
        # two collates in a row into the same static buffers (different molecules, same shapes) stay exact
        if ids == [0, 1, 2]:
            ids2 = [3, 1, 0]
            conf2 = torch.stack([conf_indices(int(mols[i].nodes["n1"].data["xyz"].shape[1]), C, gen) for i in ids2])
            spec2 = store.prepare(ids2, conf_idxs=conf2)
            if spec2.signature == spec.signature:
                store.collate_into(spec2, db)
                torch.cuda.synchronize()
                check_sections(spec2, db.ibuf.cpu().numpy(), db.fbuf.cpu().numpy(), host_reference(mols, ids2, conf2, store.energy_parts))


@pytest.mark.gpu
def test_train_step_from_store_equals_host_batch_step():
    """A training step fed by the device collate gives the same loss and the same updated weights as the step fed by
    the host-packed batch of the same molecules and conformations."""
    from grappa_b200.model import GrappaModel
    from grappa_b200.train_step import TrainStep
    import sys, os
    sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "oracle"))
    from weights import deterministic_fill_
    dev = torch.device("cuda:0")
    mols = [S.trimer_molecule(seed=s, n_confs=6) for s in range(4)]
    conf = torch.stack([torch.arange(6) for _ in range(4)])
    results = []
    for mode in ("host", "store"):
        m = GrappaModel.from_default_config()
        sd = m.state_dict()
        deterministic_fill_(sd)
        m.load_state_dict(sd)
        for mod in m.modules():
            if hasattr(mod, "p") and isinstance(getattr(mod, "p"), float):
                mod.p = 0.0
        ts = TrainStep(m, dev, use_cuda_graph=False)
        if mode == "host":
            loss = ts.step(PackedBatch(batch([set_number_confs(g, 6, conf_idxs=conf[0]) for g in mols])))
        else:
            store = MoleculeStore(mols, dev)
            spec = store.prepare([0, 1, 2, 3], conf_idxs=conf)
            st = ts._state_for(spec)
            store.collate_into(spec, st["batch"])
            loss = ts.compute_gradients(spec, load=False)
            ts.all_reduce_gradients()
            ts.optimizer_step()
        torch.cuda.synchronize()
        results.append((float(loss), ts.flat.params.clone()))
    assert results[0][0] == results[1][0]
    assert torch.equal(results[0][1], results[1][1])
