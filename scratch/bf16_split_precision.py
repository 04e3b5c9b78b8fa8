"""Precision of split-operand GEMMs through the whole parameter path (CPU emulation; development probe / DESIGN table):
every F.linear of the oracle is replaced by an emulation of  A_hi B_hi + A_hi B_lo + A_lo B_hi  with the operand planes
rounded to bf16 (2 planes = 16 mantissa bits) or to tf32 (hi) + tf32 (lo), products and sums exact (float64), and the
resulting bonded parameters are compared with the plain fp32 oracle."""
import sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/oracle")
import numpy as np, torch, torch.nn.functional as F
import grappa_oracle as O
from weights import deterministic_fill_
from grappa_b200 import synthetic as S
from grappa_b200.model import GrappaModel

def round_tf32(x):
    i = x.contiguous().view(torch.int32)
    return ((i + 0x1000) & ~0x1FFF).view(torch.float32)
def planes(x, mode):
    if mode == "bf16x2":
        hi = x.to(torch.bfloat16).float(); lo = (x - hi).to(torch.bfloat16).float(); return [hi, lo]
    if mode == "bf16x3":
        hi = x.to(torch.bfloat16).float(); mid = (x - hi).to(torch.bfloat16).float(); lo = (x - hi - mid).to(torch.bfloat16).float(); return [hi, mid, lo]
    if mode == "tf32x1":
        return [round_tf32(x)]
    if mode == "bf16x1":
        return [x.to(torch.bfloat16).float()]
    if mode == "tf32x2":
        hi = (x.contiguous().view(torch.int32) & ~0x1FFF).view(torch.float32); lo = round_tf32(x - hi); return [hi, lo]
def make_linear(mode):
    orig = F.linear
    def lin(x, w, b=None):
        if mode == "fp32" or w.shape[0] <= 8 or w.shape[1] < 32:
            return orig(x, w, b)
        xp, wp = planes(x.detach(), mode), planes(w.detach(), mode)
        acc = torch.zeros(x.shape[:-1] + (w.shape[0],), dtype=torch.float64)
        for i, a in enumerate(xp):
            for j, c in enumerate(wp):
                if i + j < len(xp):                      # drop the terms below the kept precision (lo*lo ...)
                    acc += a.double() @ c.double().t()
        y = acc.float()
        return y + b if b is not None else y
    return lin, orig

def run(g, sd, mode):
    lin, orig = make_linear(mode)
    F.linear = lin
    try:
        with torch.no_grad():
            h, p = O.model_forward(sd, g)
    finally:
        F.linear = orig
    return p

torch.set_num_threads(8)
sd = GrappaModel.from_default_config().state_dict(); deterministic_fill_(sd)
cases = {"cfg1 (22 atoms)": S.cfg1_graph(n_confs=1, seed=0), "cfg2 batch of 5": __import__("grappa_b200.graph", fromlist=["batch"]).batch([S.trimer_molecule(s, 1) for s in range(5)]),
         "cfg3 200 residues": S.make_molecule(S.template_polymer(S.alternating_sequence(200)), seed=0, n_confs=None)}
print("| case | split | bond k | bond eq | angle k | angle eq | proper k (abs/k_std) | improper k (abs/k_std) |")
for name, g in cases.items():
    ref = run(g, sd, "fp32")
    for mode in ("tf32x1", "bf16x1", "tf32x2", "bf16x2", "bf16x3"):
        p = run(g, sd, mode)
        row = []
        for lvl, key in (("n2", "k"), ("n2", "eq"), ("n3", "k"), ("n3", "eq")):
            r = ref[lvl][key]; row.append(float(((p[lvl][key] - r).abs() / r.abs()).max()))
        for lvl in ("n4", "n4_improper"):
            r = ref[lvl]["k"]; row.append(float((p[lvl]["k"] - r).abs().max() / 1.3))
        print(f"| {name} | {mode} | " + " | ".join(f"{v:.1e}" for v in row) + " |", flush=True)
