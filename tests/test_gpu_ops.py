"""Kernel-level parity: every C-ABI op (through its autograd wrapper) against a plain PyTorch fp32 CPU restatement of
the same formula, forward and backward.  Tolerance: rtol 1e-4 (BASELINE.json north_star), atol scaled to the data."""
import math

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from grappa_b200 import ops, synthetic as S
from grappa_b200.graph import batch
from grappa_b200.plan import build_plan, get_plan

DEV = "cuda"


def close(a, b, rtol=1e-4, atol=None, msg=""):
    a = a.detach().cpu().double()
    b = b.detach().cpu().double()
    assert a.shape == b.shape, f"{msg}: shape {tuple(a.shape)} vs {tuple(b.shape)}"
    if atol is None:
        atol = 1e-5 * max(float(b.abs().max()) if b.numel() else 0.0, 1e-3)
    err = (a - b).abs()
    tol = atol + rtol * b.abs()
    if not bool((err <= tol).all()):
        i = int(torch.argmax(err - tol))
        raise AssertionError(f"{msg}: max abs err {float(err.max()):.3e} (ref scale {float(b.abs().max()):.3e}); "
                             f"worst at flat index {i}: got {float(a.flatten()[i]):.7g} want {float(b.flatten()[i]):.7g}")


def hash_uniform(seed, idx):
    """Python mirror of grappa::hash_uniform (csrc/common.cuh)."""
    M = (1 << 64) - 1
    z = (seed + 0x9E3779B97F4A7C15 * (idx + 1)) & M
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M
    z = z ^ (z >> 31)
    return (z >> 40) / 16777216.0


@pytest.mark.parametrize("M,N,K", [(1, 1, 1), (22, 512, 83), (130, 511, 256), (257, 2, 256), (300, 6, 256), (64, 64, 16),
                                   (1280, 2048, 512), (1000, 512, 2048), (2100, 256, 1536), (3, 1536, 512)])
@pytest.mark.parametrize("act", [0, 1])
def test_linear_forward_backward(M, N, K, act):
    g = torch.Generator().manual_seed(M * 7 + N * 3 + K + act)
    x = torch.randn(M, K, generator=g)
    w = torch.randn(N, K, generator=g) / math.sqrt(K)
    b = torch.randn(N, generator=g) * 0.1
    r1 = torch.randn(M, N, generator=g)
    r2 = torch.randn(M, N, generator=g)
    gy = torch.randn(M, N, generator=g)
    xs = [t.clone().requires_grad_(True) for t in (x, w, b, r1, r2)]
    z = F.linear(xs[0], xs[1], xs[2])
    ref = (F.elu(z) if act else z) + xs[3] + xs[4]
    ref.backward(gy)
    ys = [t.clone().to(DEV).requires_grad_(True) for t in (x, w, b, r1, r2)]
    out = ops.linear(ys[0], ys[1], ys[2], act=act, res1=ys[3], res2=ys[4])
    out.backward(gy.to(DEV))
    close(out, ref, msg="forward")
    for name, a, r in zip(("dx", "dw", "db", "dres1", "dres2"), ys, xs):
        close(a.grad, r.grad, msg=name, atol=2e-5 * max(float(r.grad.abs().max()), 1e-3))


def test_linear_padded_output_and_no_bias():
    g = torch.Generator().manual_seed(5)
    x, w = torch.randn(50, 256, generator=g), torch.randn(511, 256, generator=g) / 16
    xg, wg = x.to(DEV).requires_grad_(True), w.to(DEV).requires_grad_(True)
    y = ops.linear(xg, wg, None, act=1, out_ld=512)
    assert y.shape == (50, 512) and float(y[:, 511].abs().max()) == 0.0
    gy = torch.randn(50, 512, generator=g)
    gy[:, 511] = 0
    y.backward(gy.to(DEV))
    xr, wr = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
    yr = F.elu(F.linear(xr, wr))
    yr.backward(gy[:, :511])
    close(y[:, :511], yr, msg="fwd")
    close(xg.grad, xr.grad, msg="dx")
    close(wg.grad, wr.grad, msg="dw")


def test_linear_strided_input_rows():
    g = torch.Generator().manual_seed(6)
    big = torch.randn(40, 512, generator=g)
    w = torch.randn(64, 511, generator=g) / 22
    y = ops.linear(big.to(DEV)[:, :511], w.to(DEV))
    close(y, F.linear(big[:, :511], w), msg="strided rows")


def test_dropout_epilogue_matches_hash_and_backward_is_consistent():
    M, N, K, p = 37, 50, 24, 0.3
    g = torch.Generator().manual_seed(1)
    x, w = torch.randn(M, K, generator=g), torch.randn(N, K, generator=g)
    res = torch.randn(M, N, generator=g)
    ops.DropoutSeeds.manual_seed(1234)
    xg = x.to(DEV).requires_grad_(True)
    y = ops.linear(xg, w.to(DEV), None, act=1, res1=res.to(DEV), drop_p=p)
    seed = ops.DropoutSeeds.base
    keep = torch.tensor([[hash_uniform(seed, m * N + n) >= p for n in range(N)] for m in range(M)])
    assert 0.55 < keep.float().mean() < 0.85
    xr = x.clone().requires_grad_(True)
    ref = F.elu(F.linear(xr, w)) * keep / (1 - p) + res
    close(y, ref, msg="dropout forward")
    gy = torch.randn(M, N, generator=g)
    y.backward(gy.to(DEV))
    ref.backward(gy)
    close(xg.grad, xr.grad, msg="dropout backward")
    # a device-side seed offset changes the mask (what CUDA-graph replays rely on)
    ops.DropoutSeeds.manual_seed(1234)
    ops.DropoutSeeds.dev = torch.tensor([77], dtype=torch.int64, device=DEV)
    try:
        y2 = ops.linear(x.to(DEV), w.to(DEV), None, act=1, res1=res.to(DEV), drop_p=p)
    finally:
        ops.DropoutSeeds.dev = None
    assert not torch.equal(y2, y.detach())
    keep2 = torch.tensor([[hash_uniform((seed + 77) & ((1 << 64) - 1), m * N + n) >= p for n in range(N)] for m in range(M)])
    close(y2, (F.elu(F.linear(x, w)) * keep2 / (1 - p) + res), msg="device seed offset")


@pytest.mark.parametrize("M,D", [(1, 256), (5, 512), (1000, 512), (333, 2048), (70, 1536), (9, 100), (600, 1024)])
def test_layernorm(M, D):
    g = torch.Generator().manual_seed(M + D)
    x = torch.randn(M, D, generator=g) * 2 + 0.5
    gamma, beta = torch.rand(D, generator=g) + 0.5, torch.randn(D, generator=g)
    gy = torch.randn(M, D, generator=g)
    r = [t.clone().requires_grad_(True) for t in (x, gamma, beta)]
    ref = F.layer_norm(r[0], (D,), r[1], r[2])
    ref.backward(gy)
    a = [t.clone().to(DEV).requires_grad_(True) for t in (x, gamma, beta)]
    out = ops.layer_norm(*a)
    out.backward(gy.to(DEV))
    close(out, ref, msg="ln fwd")
    for name, u, v in zip(("dx", "dgamma", "dbeta"), a, r):
        close(u.grad, v.grad, msg=name, atol=3e-5 * max(float(v.grad.abs().max()), 1e-3))


def test_featurize_matches_reference_concat():
    import grappa_oracle as O
    g = S.make_molecule(S.template_polymer("AFAF"), seed=3, n_confs=None, random_ring_encoding=True)
    n1 = dict(g.nodes["n1"].data)
    n1["partial_charge"] = torch.tensor(np.linspace(-2.5, 2.5, g.num_nodes("n1")), dtype=torch.float32)  # incl. clamp + bin edges
    n1["partial_charge"][:4] = torch.tensor([-2.0, 2.0, 0.0, 2.0 - 4 / 15])
    ref = O.input_features(n1)
    from grappa_b200.model import GrappaModel
    gnn = GrappaModel.from_default_config().gnn
    dev = {k: v.to(DEV) for k, v in n1.items()}
    out = ops.featurize(gnn._feature_blocks(dev), dev["partial_charge"])
    assert out.shape == ref.shape == (g.num_nodes("n1"), 83)
    assert torch.equal(out.cpu(), ref), (out.cpu() - ref).abs().max()


def _attn_ref(ft, src, dst, H):
    import grappa_oracle as O
    return O.dot_gat_conv(ft, src, dst, H)


@pytest.mark.parametrize("H,D", [(16, 32), (8, 64), (8, 16), (4, 32)])
def test_gnn_attention_forward_backward(H, D):
    g = batch([S.trimer_molecule(s, 1) for s in (0, 1, 2)])
    plan = build_plan(g).to(DEV)
    src, dst = g.edges()
    N = g.num_nodes("n1")
    gen = torch.Generator().manual_seed(H * D)
    ft = torch.randn(N, H * D, generator=gen)
    gy = torch.randn(N, H * D, generator=gen)
    fr = ft.clone().requires_grad_(True)
    ref = _attn_ref(fr, src, dst, H)
    ref.backward(gy)
    fg = ft.to(DEV).requires_grad_(True)
    out = ops.gnn_attention(fg, plan, H)
    out.backward(gy.to(DEV))
    close(out, ref, msg="attn fwd")
    close(fg.grad, fr.grad, msg="attn bwd", atol=3e-5 * float(fr.grad.abs().max()))


def test_gnn_attention_properties():
    """DotGatConv has no reference-side golden ("parity unpinned"): check the defining properties instead."""
    g = S.trimer_molecule(1, 1)
    N, H, D = g.num_nodes("n1"), 16, 32
    plan = build_plan(g).to(DEV)
    ft = torch.randn(N, H * D, generator=torch.Generator().manual_seed(0)).to(DEV)
    out = ops.gnn_attention(ft, plan, H)
    # (1) a node with a single incoming edge copies its neighbour exactly (softmax over one edge = 1)
    src, dst = g.edges()
    deg = torch.bincount(dst.long(), minlength=N)
    for v in torch.nonzero(deg == 1).flatten().tolist():
        u = int(src[dst == v][0])
        close(out[v], ft[u], msg="degree-1 copy", atol=1e-6)
    # (2) invariance to the order of the edge list
    perm = torch.randperm(g.num_edges(), generator=torch.Generator().manual_seed(1))
    g2 = S.trimer_molecule(1, 1)
    g2._edges = g2._edges[:, perm]
    out2 = ops.gnn_attention(ft, build_plan(g2).to(DEV), H)
    close(out2, out, msg="edge order invariance", atol=1e-5)
    # (3) outputs are convex combinations per head: within the min/max of the neighbours' features
    for v in (0, 5, 17):
        nb = src[dst == v].long().to(DEV)
        assert bool((out[v] <= ft[nb].max(0).values + 1e-5).all()) and bool((out[v] >= ft[nb].min(0).values - 1e-5).all())


@pytest.mark.parametrize("level,n_seq,Fdim", [("n2", 2, 512), ("n3", 3, 511), ("n4", 4, 511), ("n4_improper", 4, 511)])
def test_tuple_gather(level, n_seq, Fdim):
    g = batch([S.trimer_molecule(s, 1) for s in (0, 1)])
    plan = build_plan(g).to(DEV)
    N, T = g.num_nodes("n1"), g.num_nodes(level)
    gen = torch.Generator().manual_seed(n_seq)
    a = torch.randn(N, 512, generator=gen)
    pe = None if Fdim == 512 else torch.tensor([0., 1., 0.] if n_seq == 3 else [0., 1., 1., 0.])
    idx = g.nodes[level].data["idxs"]
    ar = a.clone().requires_grad_(True)
    ref = ar[:, :Fdim][idx]                                   # (T, n_seq, F)
    if pe is not None:
        ref = torch.cat([ref, pe.view(1, n_seq, 1).expand(T, n_seq, 1)], dim=-1)
    ref = ref.reshape(T * n_seq, 512)
    gy = torch.randn(T * n_seq, 512, generator=gen)
    ref.backward(gy)
    ag = a.to(DEV).requires_grad_(True)
    out = ops.tuple_gather(ag, plan, level, n_seq, Fdim, 512, None if pe is None else pe.to(DEV))
    out.backward(gy.to(DEV))
    assert torch.equal(out.cpu(), ref.detach())
    close(ag.grad, ar.grad, msg="gather bwd", atol=1e-5)
    if Fdim < 512:
        assert float(ag.grad[:, Fdim:].abs().max()) == 0.0


@pytest.mark.parametrize("n_seq", [2, 3, 4])
def test_tuple_attention(n_seq):
    T, H, E = 77, 8, 512
    gen = torch.Generator().manual_seed(n_seq)
    qkv = torch.randn(T * n_seq, 3 * E, generator=gen)
    gy = torch.randn(T * n_seq, E, generator=gen)
    qr = qkv.clone().requires_grad_(True)
    q, k, v = [t.view(T, n_seq, H, 64).transpose(1, 2) for t in qr.split(E, dim=1)]
    ref = F.scaled_dot_product_attention(q, k, v).transpose(1, 2).reshape(T * n_seq, E)
    ref.backward(gy)
    qg = qkv.to(DEV).requires_grad_(True)
    out = ops.tuple_attention(qg, T, n_seq, H)
    out.backward(gy.to(DEV))
    close(out, ref, msg="tuple attn fwd")
    close(qg.grad, qr.grad, msg="tuple attn bwd", atol=2e-5 * float(qr.grad.abs().max()))


def test_tuple_attention_with_the_qkv_gather_fused_is_bit_identical():
    """The QKV rows of the (atom, pe) table are gathered inside the attention kernels: same output and same table
    gradient, bit for bit, as materialising the gathered (T*n_seq, 3E) tensor first."""
    from grappa_b200.graph import batch
    from grappa_b200.plan import build_plan
    g = batch([S.trimer_molecule(seed=s, n_confs=1) for s in (0, 1, 2)])
    plan = build_plan(g).to(DEV)
    for lvl, n_seq in (("n3", 3), ("n4", 4)):
        names = ("tab_idx_" + lvl, "tab_ptr_" + lvl, "tab_tok_" + lvl)
        T, R = plan.T[lvl], 2 * plan.n_atoms
        gen = torch.Generator().manual_seed(n_seq)
        tab = torch.randn(R, 1536, generator=gen).to(DEV)
        gy = torch.randn(T * n_seq, 512, generator=gen).to(DEV)
        a = tab.clone().requires_grad_(True)
        out_a = ops.tuple_attention(ops.row_gather(a, plan, names, T * n_seq), T, n_seq, 8)
        out_a.backward(gy)
        b = tab.clone().requires_grad_(True)
        out_b = ops.tuple_attention(b, T, n_seq, 8, gather=(plan, names))
        out_b.backward(gy)
        assert torch.equal(out_a, out_b) and torch.equal(a.grad, b.grad)


@pytest.mark.parametrize("n_seq,perms", [(2, [[0, 1], [1, 0]]), (3, [[0, 1, 2], [2, 1, 0]]), (4, [[0, 1, 2, 3], [3, 1, 2, 0]])])
def test_symmetriser_permute_with_fused_layer_norm(n_seq, perms):
    """grappa_symm_ln_*: permuted copies + LayerNorm over the (n_seq * W) row in one kernel, against
    F.layer_norm of the explicitly stacked permutations (forward, dX, dgamma, dbeta)."""
    T, W = 333, 512
    gen = torch.Generator().manual_seed(n_seq)
    x = torch.randn(T * n_seq, W, generator=gen) * 1.7 + 0.3
    gamma, beta = 1 + 0.1 * torch.randn(n_seq * W, generator=gen), 0.1 * torch.randn(n_seq * W, generator=gen)
    gy = torch.randn(len(perms) * T, n_seq * W, generator=gen)
    xr, gr, br = [t.clone().double().requires_grad_(True) for t in (x, gamma, beta)]
    x3 = xr.view(T, n_seq, W)
    stacked = torch.cat([x3[:, p, :].reshape(T, n_seq * W) for p in perms], dim=0)
    ref = F.layer_norm(stacked, (n_seq * W,), gr, br)
    ref.backward(gy.double())
    xg, gg, bg = [t.clone().to(DEV).requires_grad_(True) for t in (x, gamma, beta)]
    out = ops.symm_layer_norm(xg, gg, bg, T, n_seq, perms)
    out.backward(gy.to(DEV))
    close(out, ref, msg="symm ln fwd")
    close(xg.grad, xr.grad, msg="symm ln dx", atol=2e-5 * float(xr.grad.abs().max()))
    close(gg.grad, gr.grad, msg="symm ln dgamma", atol=2e-5 * float(gr.grad.abs().max()))
    close(bg.grad, br.grad, msg="symm ln dbeta", atol=2e-5 * float(br.grad.abs().max()))
    # identical to the two-kernel path up to the summation order of the statistics
    two = ops.layer_norm(ops.symm_permute(xg.detach(), T, n_seq, perms), gg.detach(), bg.detach())
    assert float((two - out.detach()).abs().max()) < 2e-5


def test_tuple_attention_matches_nn_multihead_attention():
    """The composition in_proj GEMM -> tuple attention -> out_proj GEMM equals nn.MultiheadAttention (seq-first)."""
    T, S_, E = 31, 3, 512
    mha = torch.nn.MultiheadAttention(E, 8, dropout=0)
    x = torch.randn(S_, T, E, generator=torch.Generator().manual_seed(0))
    ref, _ = mha(x, x, x, need_weights=False)
    xt = x.transpose(0, 1).reshape(T * S_, E).to(DEV)
    qkv = ops.linear(xt, mha.in_proj_weight.detach().to(DEV), mha.in_proj_bias.detach().to(DEV))
    o = ops.tuple_attention(qkv, T, S_, 8)
    y = ops.linear(o, mha.out_proj.weight.detach().to(DEV), mha.out_proj.bias.detach().to(DEV))
    close(y.view(T, S_, E).transpose(0, 1), ref, msg="mha composition")


def test_token_residual_and_symm_permute():
    T, n_seq, W, Fd = 19, 4, 512, 511
    gen = torch.Generator().manual_seed(2)
    x3, x = torch.randn(T * n_seq, W, generator=gen), torch.randn(T * n_seq, W, generator=gen)
    a, b = x3.clone().requires_grad_(True), x.clone().requires_grad_(True)
    ref = a + b + torch.cat([b[:, :Fd], torch.zeros(T * n_seq, W - Fd)], dim=1)
    gy = torch.randn(T * n_seq, W, generator=gen)
    ref.backward(gy)
    ag, bg = x3.to(DEV).requires_grad_(True), x.to(DEV).requires_grad_(True)
    out = ops.token_residual(ag, bg, Fd)
    out.backward(gy.to(DEV))
    close(out, ref, msg="token residual")
    close(ag.grad, a.grad, msg="d x3")
    close(bg.grad, b.grad, msg="d x")
    for perms in ([[0, 1, 2, 3], [3, 2, 1, 0]], [[0, 1, 2, 3], [3, 1, 2, 0]]):
        xr = x.clone().requires_grad_(True)
        xs = xr.view(T, n_seq, W).transpose(0, 1)                      # reference layout (n_seq, T, W)
        ref = torch.stack([xs[p] for p in perms], 0).transpose(1, 2).contiguous().view(2 * T, n_seq * W)
        gy = torch.randn(2 * T, n_seq * W, generator=gen)
        ref.backward(gy)
        xg = x.to(DEV).requires_grad_(True)
        out = ops.symm_permute(xg, T, n_seq, perms)
        out.backward(gy.to(DEV))
        assert torch.equal(out.cpu(), ref.detach())
        close(xg.grad, xr.grad, msg="permute bwd", atol=1e-5)


def test_readout_finalize_all_kinds():
    import grappa_oracle as O
    T = 53
    gen = torch.Generator().manual_seed(3)
    pref = [1.0, 1.0]
    # bond / angle
    for kind, consts in ((0, [6.3, 0.195, 0.0, 4.7, 161.2, 0.0, 0.0]), (1, [0.0292, math.pi, 0.0, 3.97, 26.6, 0.0, 0.0])):
        s = torch.randn(2 * T, 2, generator=gen)
        sr = s.clone().requires_grad_(True)
        c = sr.view(2, T, 2).sum(0)
        if kind == 0:
            eq = consts[1] * (F.elu(consts[0] + c[:, 0] - 1) + 1) + consts[2]
        else:
            eq = consts[1] * torch.sigmoid(consts[0] * c[:, 0])
        k = consts[4] * (F.elu(consts[3] + c[:, 1] - 1) + 1) + consts[5]
        g1, g2 = torch.randn(T, generator=gen), torch.randn(T, generator=gen)
        (eq * g1 + k * g2).sum().backward()
        sg = s.to(DEV).requires_grad_(True)
        eq_g, k_g = ops.readout_finalize(sg, kind, T, pref, consts)
        (eq_g * g1.to(DEV) + k_g * g2.to(DEV)).sum().backward()
        close(eq_g, eq, msg=f"kind {kind} eq")
        close(k_g, k, msg=f"kind {kind} k")
        close(sg.grad, sr.grad, msg=f"kind {kind} grad")
    # gated torsions: proper (3 of 3 columns) and improper (only n = 2), with entries on both sides of the cutoff
    for n_out, n_per, col0 in ((6, 3, 0), (2, 2, 1)):
        nh = n_out // 2
        kstd = torch.rand(nh, generator=gen) + 0.5
        s = torch.randn(2 * T, n_out, generator=gen)
        s[:6, :nh] = 1e-5
        sr = s.clone().requires_grad_(True)
        c = sr.view(2, T, n_out).sum(0)
        k = O.hard_cutoff(c[:, :nh] * torch.sigmoid(c[:, nh:]) * kstd, 1e-4)
        full = torch.zeros(T, n_per)
        full = torch.cat([torch.zeros(T, col0), k, torch.zeros(T, n_per - col0 - nh)], dim=1)
        gk = torch.randn(T, n_per, generator=gen)
        (full * gk).sum().backward()
        sg = s.to(DEV).requires_grad_(True)
        k_g = ops.readout_finalize(sg, 2, T, pref, [0, 0, 0, 0, 0, 0, 1e-4], kstd=kstd.to(DEV), kmean=torch.zeros(nh, device=DEV),
                                   n_per=n_per, col0=col0)
        (k_g * gk.to(DEV)).sum().backward()
        close(k_g, full, msg="torsion k", atol=1e-7)
        assert bool(((k_g == 0).cpu() == (full == 0)).all()), "cutoff decisions differ"
        close(sg.grad, sr.grad, msg="torsion grad", atol=1e-6)


def test_empty_inputs_are_handled():
    x = torch.zeros((0, 512), device=DEV)
    w = torch.randn(256, 512, device=DEV)
    assert ops.linear(x, w).shape == (0, 256)
    assert ops.layer_norm(x, torch.ones(512, device=DEV), torch.zeros(512, device=DEV)).shape == (0, 512)
    assert ops.tuple_attention(torch.zeros((0, 1536), device=DEV), 0, 3, 8).shape == (0, 512)


def test_pe_table_and_row_gather_ops():
    """Token de-duplication helpers: the (atom, pe) table and the row gather with its pull-style backward."""
    from grappa_b200.plan import PE_PATTERN
    g = batch([S.trimer_molecule(s, 2) for s in (0, 1)]).to(DEV)
    plan = get_plan(g)
    N = g.num_nodes("n1")
    torch.manual_seed(0)
    for level, n_seq in (("n3", 3), ("n4", 4), ("n4_improper", 4)):
        T = plan.T[level]
        P = torch.randn(N, 512, device=DEV, requires_grad=True)
        U = ops.pe_table(P, 511, 512)
        ref_U = torch.stack([torch.cat([P[:, :511], torch.full((N, 1), float(pe), device=DEV)], 1) for pe in (0, 1)], 1).reshape(2 * N, 512)
        assert torch.equal(U, ref_U)
        names = ("tab_idx_" + level, "tab_ptr_" + level, "tab_tok_" + level)
        X = ops.row_gather(U, plan, names, T * n_seq)
        idx = g.nodes[level].data["idxs"].long()
        pe = torch.tensor(PE_PATTERN[level], device=DEV)
        rows = (2 * idx + pe[None, :]).reshape(-1)
        assert torch.equal(plan.view(names[0]).long(), rows)
        assert torch.equal(X, ref_U[rows])
        # the same tokens as the direct gather of the writers
        direct = ops.tuple_gather(P, plan, level, n_seq, 511, 512, pe.float())
        assert torch.equal(X, direct)
        w = torch.randn_like(X)
        (X * w).sum().backward()
        Pr = P.detach().clone().requires_grad_(True)
        Ur = torch.stack([torch.cat([Pr[:, :511], torch.full((N, 1), float(q), device=DEV)], 1) for q in (0, 1)], 1).reshape(2 * N, 512)
        (Ur[rows] * w).sum().backward()
        torch.testing.assert_close(P.grad, Pr.grad, rtol=1e-5, atol=1e-5)
    # wide rows (the QKV table) through the bond-level incidence lists, i.e. no PE column
    tab = torch.randn(N, 1536, device=DEV, requires_grad=True)
    names = ("idx_n2", "inc_ptr_n2", "inc_tok_n2")
    X = ops.row_gather(tab, plan, names, plan.T["n2"] * 2)
    rows = g.nodes["n2"].data["idxs"].long().reshape(-1)
    assert torch.equal(X, tab[rows])
    w = torch.randn_like(X)
    (X * w).sum().backward()
    ref = torch.zeros_like(tab).index_add_(0, rows, w)
    torch.testing.assert_close(tab.grad, ref, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("M,H,act2", [(300, 256, 0), (1280, 2048, 1), (77, 100, 0), (4096, 2048, 0)])
def test_elu_mlp_fused_activation_gradient(M, H, act2, monkeypatch):
    """Linear-ELU-Linear with ELU' fused into the second layer's dX GEMM epilogue (grappa_gemm_actgrad) against the
    unfused path and against PyTorch: interior and edge tiles, split-K and FFMA fallbacks included via the shapes."""
    torch.manual_seed(M + H)
    D = 512 if H >= 256 else 52
    x = torch.randn(M, D, device=DEV)
    w1, b1 = torch.randn(H, D, device=DEV) / math.sqrt(D), torch.randn(H, device=DEV) * 0.1
    w2, b2 = torch.randn(D, H, device=DEV) / math.sqrt(H), torch.randn(D, device=DEV) * 0.1
    gout = torch.randn(M, D, device=DEV)
    res = {}
    for mode in ("fused", "unfused", "torch"):
        leaves = [t.clone().requires_grad_(True) for t in (x, w1, b1, w2, b2)]
        if mode == "torch":
            z = F.elu(leaves[0] @ leaves[1].T + leaves[2])
            y = z @ leaves[3].T + leaves[4]
            y = F.elu(y) if act2 else y
            y = y + leaves[0]
        else:
            monkeypatch.setattr(ops, "FUSE_ELU_GRAD", mode == "fused")
            y = ops.elu_mlp(leaves[0], leaves[1], leaves[2], leaves[3], leaves[4], act2=act2, res1=leaves[0])
        y.backward(gout)
        res[mode] = [y.detach()] + [t.grad for t in leaves]
    for a, b, c in zip(res["fused"], res["unfused"], res["torch"]):
        scale = float(c.abs().max()) + 1e-12
        assert float((a - b).abs().max()) / scale < 2e-6
        assert float((a - c).abs().max()) / scale < 2e-5


# relative-to-max error bars of one dense layer per GEMM math mode: 1 = 3xTF32 (~21 mantissa bits per operand),
# 2 = bf16x3 for forward / dX (16 mantissa bits per operand: measured 4e-6 .. 6e-6) with 3xTF32 weight gradients
GEMM_TOL = {1: 3e-6, 2: 1.2e-5}


@pytest.mark.parametrize("mode", [1, 2])
@pytest.mark.parametrize("M,N,K", [(1280, 2048, 512), (6528, 512, 512), (12160, 512, 256), (2000, 1300, 96), (19072, 128, 64)])
def test_linear_multi_wave_grids(M, N, K, mode, monkeypatch):
    """Tile grids larger than one wave of the 148 SMs (160 ... 1520 tiles): all epilogue variants and a ragged edge,
    forward and backward, run-to-run determinism; both tensor-core math modes."""
    monkeypatch.setattr(ops, "MATH_MODE", mode)
    tol = GEMM_TOL[mode]
    torch.manual_seed(M + N + K)
    x = torch.randn(M, K, device=DEV)
    w = torch.randn(N, K, device=DEV) / math.sqrt(K)
    b = torch.randn(N, device=DEV) * 0.1
    r1, r2 = torch.randn(M, N, device=DEV), torch.randn(M, N, device=DEV)
    ref = x.double() @ w.double().T + b.double()
    for act, nres in ((0, 0), (1, 1), (0, 2)):
        y = ops.linear(x, w, b, act=act, res1=r1 if nres >= 1 else None, res2=r2 if nres >= 2 else None)
        want = F.elu(ref) if act else ref
        want = want + (r1.double() if nres >= 1 else 0) + (r2.double() if nres >= 2 else 0)
        err = float((y.double() - want).abs().max()) / float(want.abs().max())
        assert err < tol, (act, nres, err)
    xg, wg = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
    gout = torch.randn(M, N, device=DEV)
    ops.linear(xg, wg, b).backward(gout)
    dx = gout.double() @ w.double()
    dw = gout.double().T @ x.double()
    assert float((xg.grad.double() - dx).abs().max()) / float(dx.abs().max()) < tol
    assert float((wg.grad.double() - dw).abs().max()) / float(dw.abs().max()) < 3e-6        # weight gradients: 3xTF32 in both modes
    # run-to-run determinism
    assert torch.equal(ops.linear(x, w, b), ops.linear(x, w, b))


@pytest.mark.parametrize("M,N,K", [(300, 512, 256), (1280, 2048, 512), (2000, 130, 96)])
def test_presplit_weight_operand_is_bit_identical(M, N, K, monkeypatch):
    """grappa_gemm_ex with B_lo from grappa_split_lo (weights split once instead of in every CTA and k-block) feeds the
    tensor core exactly the operands the in-kernel split produces: forward and dX are bit-identical."""
    torch.manual_seed(M + N)
    x = torch.randn(M, K, device=DEV)
    w = (torch.randn(N, K, device=DEV) / math.sqrt(K)).requires_grad_(True)
    b = torch.randn(N, device=DEV) * 0.1
    gout = torch.randn(M, N, device=DEV)
    lo = ops.split_lo(w.detach())
    hi = (w.detach().view(torch.int32) & -8192).view(torch.float32)
    rec = hi.double() + lo.double()
    assert float((rec - w.detach().double()).abs().max()) <= float(w.detach().abs().max()) * 2.0 ** -21
    outs = []
    for use in (False, True):
        monkeypatch.setattr(ops, "WEIGHT_LO", {id(w): lo} if use else {})
        xg = x.clone().requires_grad_(True)
        y = ops.linear(xg, w, b, act=ops.ACT_ELU)
        y.backward(gout)
        outs.append((y.detach().clone(), xg.grad.clone()))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])


@pytest.mark.parametrize("mode", [1, 2])
@pytest.mark.parametrize("M,N_out", [(6080, 6), (4352, 2), (100, 1), (20000, 8), (63, 2)])
def test_skinny_output_layer_forward_and_backward(M, N_out, mode, monkeypatch):
    """Last symmetriser layer (256 -> 2 / 6): forward, dX (with the fused ELU') and dW run on the streaming kernels of
    gemm_skinny.cu (one matrix dimension <= 8); against float64 PyTorch."""
    monkeypatch.setattr(ops, "MATH_MODE", mode)            # the 256 -> 256 layer in front runs on the tensor cores
    torch.manual_seed(M + N_out)
    x = torch.randn(M, 256, device=DEV)
    w1, b1 = torch.randn(256, 256, device=DEV) / 16, torch.randn(256, device=DEV) * 0.1
    w2, b2 = torch.randn(N_out, 256, device=DEV) / 16, torch.randn(N_out, device=DEV) * 0.1
    gout = torch.randn(M, N_out, device=DEV)
    leaves = [t.clone().requires_grad_(True) for t in (x, w1, b1, w2, b2)]
    y = ops.elu_mlp(*leaves)
    y.backward(gout)
    ref = [t.double().clone().requires_grad_(True) for t in (x, w1, b1, w2, b2)]
    yr = F.elu(ref[0] @ ref[1].T + ref[2]) @ ref[3].T + ref[4]
    yr.backward(gout.double())
    for a, b, name in zip([y] + [t.grad for t in leaves], [yr] + [t.grad for t in ref], ("y", "dx", "dw1", "db1", "dw2", "db2")):
        scale = float(b.abs().max()) + 1e-30
        assert float((a.double() - b).abs().max()) / scale < (5e-6 if mode == 1 else 2e-5), name
    y2 = ops.elu_mlp(*[t.detach() for t in leaves])
    assert torch.equal(y.detach(), y2)                      # deterministic
