"""
TEST INFRASTRUCTURE.  Converts the reference tree's one binary fixture,
    /root/reference/data_old/dgl_datasets/small_example/graphs.bin   (12 small molecules, 40-50 conformations each, QM
    energies / forces and the nonbonded reference contributions; data_old/dataset_tags.csv:2)
into tests/golden/small_example.npz with this repo's DGL-free reader (grappa_b200/dgl_bin.py), keeping the fields the hot
path and the loss read:

    python oracle/make_real_fixture.py

/root/reference does not travel to the GPU box; the npz does.  energy_ref / gradient_ref follow the reference's
convention: QM minus the nonbonded reference contribution (data/dataset.py: `reference_ff_nonbonded` is subtracted so
that the model only has to fit the bonded part).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from grappa_b200.dgl_bin import load_graphs  # noqa: E402

SRC = "/root/reference/data_old/dgl_datasets/small_example/graphs.bin"
OUT = os.path.join(ROOT, "tests", "golden", "small_example.npz")
N1_FIELDS = ("atomic_number", "ring_encoding", "partial_charge", "degree", "xyz", "gradient_qm", "gradient_reference_ff_nonbonded")
G_FIELDS = ("energy_qm", "energy_reference_ff_nonbonded")


def main():
    graphs = load_graphs(SRC)
    out = {"n_graphs": np.int64(len(graphs))}
    for i, g in enumerate(graphs):
        for k in N1_FIELDS:
            out[f"g{i}_n1_{k}"] = g["n1"][k]
        for k in G_FIELDS:
            out[f"g{i}_g_{k}"] = g["g"][k]
        for lvl, n_seq in (("n2", 2), ("n3", 3), ("n4", 4), ("n4_improper", 4)):
            out[f"g{i}_{lvl}_idxs"] = np.asarray(g[lvl]["idxs"], np.int64).reshape(-1, n_seq)
    np.savez_compressed(OUT, **out)
    print(f"wrote {OUT}: {os.path.getsize(OUT) / 1e3:.1f} kB, {len(graphs)} graphs")


if __name__ == "__main__":
    main()
