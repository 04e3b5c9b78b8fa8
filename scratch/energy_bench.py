"""cfg-4 energy sweep timing: python scratch/energy_bench.py [n_mol] [n_conf] [parts,parts,...]  (development probe)"""
import os, sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/oracle")
import torch
from grappa_b200 import synthetic as S
from grappa_b200.energy import Energy
n_mol = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
n_conf = int(sys.argv[2]) if len(sys.argv) > 2 else 500
parts = [int(x) for x in sys.argv[3].split(",")] if len(sys.argv) > 3 else [int(os.environ.get("GRAPPA_ENERGY_PARTS", 4))]
g = S.cfg4_batch(n_mol, n_conf).to("cuda")
gen = torch.Generator().manual_seed(0)
T = {l: g.num_nodes(l) for l in ("n2", "n3", "n4", "n4_improper")}
g.nodes["n2"].data["k"] = (500 + 300 * torch.rand(T["n2"], generator=gen)).cuda(); g.nodes["n2"].data["eq"] = (1 + .5 * torch.rand(T["n2"], generator=gen)).cuda()
g.nodes["n3"].data["k"] = (60 + 80 * torch.rand(T["n3"], generator=gen)).cuda(); g.nodes["n3"].data["eq"] = (1.5 + .8 * torch.rand(T["n3"], generator=gen)).cuda()
g.nodes["n4"].data["k"] = torch.randn(T["n4"], 3, generator=gen).cuda(); g.nodes["n4_improper"].data["k"] = torch.randn(T["n4_improper"], 2, generator=gen).cuda()
en = Energy(materialize_tuple_terms=False)
for w in parts:
    os.environ["GRAPPA_ENERGY_PARTS"] = str(w)
    if hasattr(g, "_plan_cache"):
        g._plan_cache.clear()
    with torch.no_grad():
        for _ in range(3): en(g)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            for _ in range(10): en(g)
        graph.replay(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); graph.replay(); b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 10
    units = n_mol * n_conf
    print(f"energy {n_mol}x{n_conf} parts={w}: {ms:.3f} ms, {units/ms/1e3:.1f} M mol-conf/s, {units*980/ms/1e6:.1f} GB/s algorithmic, "
          f"checksum {float(g.nodes['g'].data['energy'].double().sum()):.6e} {float(g.nodes['n1'].data['gradient'].double().abs().sum()):.6e}", flush=True)
