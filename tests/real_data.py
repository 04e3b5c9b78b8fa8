"""Helper for the tests: the reference tree's real molecules (tests/golden/small_example.npz, made by
oracle/make_real_fixture.py from data_old/dgl_datasets/small_example/graphs.bin) as HeteroGraphs."""
import os

import numpy as np
import torch

from grappa_b200.dgl_bin import to_heterograph

LEVELS = ("n2", "n3", "n4", "n4_improper")


def load_real_molecules(golden_dir, n_confs=None):
    z = np.load(os.path.join(golden_dir, "small_example.npz"))
    mols = []
    for i in range(int(z["n_graphs"])):
        xyz = z[f"g{i}_n1_xyz"]
        c = xyz.shape[1] if n_confs is None else min(n_confs, xyz.shape[1])
        n1 = {k: z[f"g{i}_n1_{k}"] for k in ("atomic_number", "ring_encoding", "partial_charge", "degree")}
        n1["xyz"] = xyz[:, :c]
        # the bonded part that the model has to reproduce: QM minus the nonbonded reference contribution
        n1["gradient_ref"] = (z[f"g{i}_n1_gradient_qm"] - z[f"g{i}_n1_gradient_reference_ff_nonbonded"])[:, :c]
        e = (z[f"g{i}_g_energy_qm"].astype(np.float64) - z[f"g{i}_g_energy_reference_ff_nonbonded"])[:, :c]
        e = e - e.mean(axis=1, keepdims=True)            # per-molecule centring, as the reference's data pipeline stores it
        data = {"n1": n1, "g": {"energy_ref": e.astype(np.float32), "is_dummy": np.zeros((1, c), np.float32)}}
        for lvl in LEVELS:
            data[lvl] = {"idxs": z[f"g{i}_{lvl}_idxs"]}
        g = to_heterograph(data)
        for lvl in LEVELS:                       # levels without tuples still exist as (empty) node types
            assert g.num_nodes(lvl) == len(z[f"g{i}_{lvl}_idxs"])
        mols.append(g)
    return mols
