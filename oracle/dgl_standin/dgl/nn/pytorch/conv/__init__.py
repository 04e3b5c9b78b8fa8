"""Stand-in for dgl.nn.pytorch.conv: only DotGatConv carries semantics (see package docstring)."""
import math
import torch


class DotGatConv(torch.nn.Module):
    """
    Restates DGL 2.4.0 `DotGatConv` (python/dgl/nn/pytorch/conv/dotgatconv.py): one bias-free Linear `fc`
    shared by source and destination, e_uv = <ft_u, ft_v> / sqrt(out_feats), softmax over the incoming
    edges of v (per head), rst_v = sum_u a_uv * ft_u.  Output (N, num_heads, out_feats).
    """
    def __init__(self, in_feats, out_feats, num_heads, allow_zero_in_degree=False):
        super().__init__()
        self._in_feats, self._out_feats, self._num_heads = in_feats, out_feats, num_heads
        self.fc = torch.nn.Linear(in_feats, out_feats * num_heads, bias=False)

    def forward(self, graph, feat, get_attention=False):
        src, dst = graph.edges()
        src, dst = src.long(), dst.long()
        n = feat.shape[0]
        ft = self.fc(feat).view(n, self._num_heads, self._out_feats)
        e = (ft[src] * ft[dst]).sum(-1) / math.sqrt(self._out_feats)          # (E, H)
        emax = torch.full((n, self._num_heads), -float("inf"), dtype=e.dtype, device=e.device)
        emax = emax.scatter_reduce(0, dst[:, None].expand_as(e), e, reduce="amax", include_self=True)
        ex = torch.exp(e - emax[dst])
        den = torch.zeros((n, self._num_heads), dtype=e.dtype, device=e.device).index_add(0, dst, ex)
        a = ex / den[dst]
        out = torch.zeros_like(ft).index_add(0, dst, a.unsqueeze(-1) * ft[src])
        return out


class _Unsupported(torch.nn.Module):
    def __init__(self, *a, **kw):
        super().__init__()
        raise NotImplementedError("dgl stand-in: only DotGatConv is implemented")


class GATConv(_Unsupported):
    pass


class GATv2Conv(_Unsupported):
    pass


class SAGEConv(_Unsupported):
    pass
