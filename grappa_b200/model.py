"""
`GrappaModel`: drop-in for `grappa.models.grappa.GrappaModel` (models/grappa.py:7-189) on the B200 kernels.

Same constructor keywords, same `forward(g) -> g` contract on a DGL-style heterograph, same `field_of_view`, and the
same state-dict layout (296 entries at the default config, including the `gnn.blocks.*` aliases, DGL's
`graph_module.fc.weight`, `attn.in_proj_weight`, the statistics buffers and the unused `grappa_transformer.norm`),
so reference checkpoints load with `load_state_dict(strict=True)`.  torch.nn modules are used only as parameter
holders / initialisers; every arithmetic step of `forward` is a kernel of libgrappa_b200.so via grappa_b200.ops.

Module names mirror the reference files they replace:
  GrappaGNN / ResidualAttentionBlock   models/graph_attention.py:11-196, :201-323
  WriteParameters and the four writers  models/interaction_parameters.py:10-615
  SymmetrisedTransformer / GrappaTransformer / Symmetriser   models/perm_equiv_transformer.py:13-337
  DottedAttWithMLP / FeedForwardLayer   models/network_utils.py:5-136
  ToPositive / ToRange                  models/final_layer.py:11-97
"""
import copy
import os
import math
from typing import Dict, List, Union

import torch
from torch import nn

from . import ops
from .plan import PE_PATTERN, get_plan

MAX_ELEMENT = 53            # constants.py:38
CHARGE_ENCODING_DIM = 16    # graph_attention.py:53


def get_default_statistics():
    """Constants of utils/graph_utils.py:309-318 (parameter statistics of a peptide dataset)."""
    return {
        "mean": {"n2_k": torch.Tensor([763.2819]), "n2_eq": torch.Tensor([1.2353]), "n3_k": torch.Tensor([105.6576]),
                 "n3_eq": torch.Tensor([1.9750]),
                 "n4_k": torch.Tensor([1.5617e-01, -5.8312e-01, 7.0820e-02, -6.3840e-04, 4.7139e-04, -4.1655e-04]),
                 "n4_improper_k": torch.Tensor([0.0000, -2.3933, 0.0000])},
        "std": {"n2_k": torch.Tensor([161.2278]), "n2_eq": torch.Tensor([0.1953]), "n3_k": torch.Tensor([26.5965]),
                "n3_eq": torch.Tensor([0.0917]), "n4_k": torch.Tensor([0.4977, 1.2465, 0.1466, 0.0192, 0.0075, 0.0066]),
                "n4_improper_k": torch.Tensor([0.0000, 4.0571, 0.0000])}}


DEFAULT_CONFIG = dict(   # configs/model/default.yaml
    graph_node_features=256, in_feats=None, in_feat_name=["atomic_number", "ring_encoding", "partial_charge", "degree"],
    in_feat_dims={}, gnn_width=512, gnn_attentional_layers=4, gnn_convolutions=0, gnn_attention_heads=16,
    gnn_dropout_attention=0.3, gnn_dropout_initial=0.0, gnn_dropout_conv=0.0, gnn_dropout_final=0.1,
    symmetric_transformer_dropout=0.5, symmetric_transformer_depth=1, symmetric_transformer_n_heads=8,
    symmetric_transformer_width=512, symmetriser_depth=4, symmetriser_width=256, n_periodicity_proper=3,
    n_periodicity_improper=2, gated_torsion=True, positional_encoding=True, layer_norm=True, self_interaction=True,
    learnable_statistics=False, torsion_cutoff=1.0e-4, harmonic_gate=False, only_n2_improper=True, stat_scaling=True,
    shifted_elu=True)


PARALLEL_WRITERS = True
_WRITER_STREAMS = {}


# Token de-duplication (see csrc/readout.cu): the first LayerNorm + QKV projection of a writer's transformer run once per
# distinct (atom, pe) row instead of once per token whenever there are at least DEDUP_RATIO times more tokens than rows.
DEDUP_TOKENS = True
DEDUP_RATIO = 1.5
FUSE_SYMM_LN = os.environ.get("GRAPPA_FUSE_SYMM_LN", "1") != "0"     # training: permute + first LayerNorm of the symmetriser in one kernel
FOLD_SYMMETRISER = True       # inference: the symmetriser's permutations folded into weight views (Symmetriser._forward_folded)
WRITER_STREAM_PRIORITIES = (-1, -1, -2, -1)


def _writer_streams(device):
    device = ops.norm_device(device)
    st = _WRITER_STREAMS.get(device)
    if st is None:
        # order: bond, angle, proper, improper.  Priorities (0 = lowest): the writers' chains sit above the side streams that
        # carry parameter-gradient work (priority 0), the proper-torsion writer (most tokens, longest chain) above the
        # rest.  Measured on the cfg-2 step: 4.35 -> 4.11 ms.
        st = [torch.cuda.Stream(device=device, priority=p) for p in WRITER_STREAM_PRIORITIES]
        ops.THROUGHPUT_STREAMS.update(x.cuda_stream for x in st)
        _WRITER_STREAMS[device] = st
    return st


def _unsupported(what: str):
    raise NotImplementedError(f"grappa_b200: {what} is outside the accelerated hot path (default-config family only)")


# ------------------------------------------------------------------------------------------------------- GNN ----

class _DotGatParams(nn.Module):
    """Parameter holder with DGL's DotGatConv naming: one bias-free Linear `fc` shared by q, k and v."""
    def __init__(self, in_feats, out_feats, num_heads):
        super().__init__()
        self.fc = nn.Linear(in_feats, out_feats * num_heads, bias=False)


class ResidualAttentionBlock(nn.Module):
    def __init__(self, in_feats, out_feats=None, num_heads=8, self_interaction=True, layer_norm=True,
                 dropout_behind_mha=0.1, dropout_behind_self_interaction=0.1, skip_connection=True):
        super().__init__()
        out_feats = in_feats if out_feats is None else out_feats
        if out_feats != in_feats or not layer_norm or not self_interaction or not skip_connection:
            _unsupported("attention blocks with out_feats != in_feats, without layer norm, self interaction or skip")
        self.in_feats, self.out_feats, self.num_heads = in_feats, out_feats, num_heads
        self.feats_per_head = out_feats // num_heads
        self.graph_module = _DotGatParams(in_feats, out_feats // num_heads, num_heads)
        self.dropout1 = nn.Dropout(p=dropout_behind_mha)
        self.dropout2 = nn.Dropout(p=dropout_behind_self_interaction)
        self.layer_norm = nn.LayerNorm(in_feats)
        self.activation = nn.ELU()
        self.head_reducer = nn.Linear(num_heads * self.feats_per_head, out_feats)
        self.interaction_norm = nn.LayerNorm(out_feats)
        self.self_interaction = nn.Sequential(nn.Linear(out_feats, 4 * out_feats), self.activation,
                                              nn.Linear(4 * out_feats, out_feats), self.activation)

    def forward(self, plan, h, pre_ffn=None):
        """graph_attention.py:289-323.  `pre_ffn` is called right before the self-interaction weights are first read."""
        p1 = self.dropout1.p if self.training else 0.0
        p2 = self.dropout2.p if self.training else 0.0
        h = ops.layer_norm(h, self.layer_norm.weight, self.layer_norm.bias)
        # h feeds fc and the skip of head_reducer: the skip's gradient is added in fc's dX GEMM epilogue (ops.LinearFn)
        holder = ops.skip_grad_holder(h)
        ft = ops.linear(h, self.graph_module.fc.weight, dx_add=holder)
        att = ops.gnn_attention(ft, plan, self.num_heads)
        h = ops.linear(att, self.head_reducer.weight, self.head_reducer.bias, res1=h, drop_p=p1, res1_grad_to=holder)
        h = ops.layer_norm(h, self.interaction_norm.weight, self.interaction_norm.bias)
        lin1, lin2 = self.self_interaction[0], self.self_interaction[2]
        if pre_ffn is not None:
            pre_ffn()
        return ops.elu_mlp(h, lin1.weight, lin1.bias, lin2.weight, lin2.bias, act2=ops.ACT_ELU, res1=h, drop_p=p2)


class ChargeEncoding(nn.Module):
    """graph_attention.py:434-464; evaluated inside the featurize kernel (bins fixed at 15 over [-2, 2])."""
    def __init__(self, min_value=-2, max_value=2, n_bins=16):
        super().__init__()
        self.min_value, self.max_value, self.n_bins = min_value, max_value, n_bins


class GrappaGNN(nn.Module):
    def __init__(self, out_feats=512, in_feats=None, node_feats=None, n_conv=3, n_att=3, n_heads=8,
                 in_feat_name=("atomic_number", "ring_encoding", "partial_charge"), in_feat_dims={}, conv_dropout=0.,
                 attention_dropout=0., final_dropout=0., initial_dropout=0., layer_norm=True, self_interaction=True,
                 charge_encoding=True):
        super().__init__()
        if n_conv != 0:
            _unsupported("gnn_convolutions > 0 (SAGEConv blocks)")
        if not charge_encoding:
            _unsupported("charge_encoding=False")
        in_feat_name = list(in_feat_name) if isinstance(in_feat_name, (list, tuple)) else [in_feat_name]
        default_dims = {"atomic_number": MAX_ELEMENT, "ring_encoding": 7, "partial_charge": 1, "sp_hybridization": 6,
                        "mass": 2, "degree": 6, "is_radical": 1, "charge_model": 4}
        default_dims.update(in_feat_dims)
        self.in_feat_dims = default_dims
        if in_feats is None:
            in_feats = sum(default_dims[f] for f in in_feat_name)
        node_feats = out_feats if node_feats is None else node_feats
        self.in_feat_name = in_feat_name
        self.in_feats = in_feats + CHARGE_ENCODING_DIM
        self.charge_encoding = True
        self.charge_encoder = ChargeEncoding(n_bins=CHARGE_ENCODING_DIM - 1, min_value=-2, max_value=2)
        self.initial_dropout = nn.Dropout(initial_dropout)
        self.final_dropout = nn.Dropout(final_dropout)
        self.pre_dense = nn.Sequential(nn.Linear(self.in_feats, node_feats), nn.ELU())
        self.no_convs = (n_conv + n_att) == 0
        if not self.no_convs:
            self.conv_blocks = nn.ModuleList([])
            self.att_blocks = nn.ModuleList([
                ResidualAttentionBlock(in_feats=node_feats, out_feats=node_feats, num_heads=n_heads,
                                       self_interaction=self_interaction, dropout_behind_mha=attention_dropout,
                                       dropout_behind_self_interaction=attention_dropout, skip_connection=True,
                                       layer_norm=layer_norm) for _ in range(n_att)])
            self.post_dense = nn.Sequential(nn.Linear(node_feats, out_feats))
            self.blocks = self.conv_blocks + self.att_blocks
        else:
            self.post_dense = nn.Sequential(nn.Linear(node_feats, out_feats))

    def _feature_blocks(self, n1):
        """graph_attention.py:156-171: sorted features first, then charge_model / atomic_number."""
        blocks = []
        for feat in sorted(self.in_feat_name):
            if feat in ("charge_model", "atomic_number"):
                continue
            t = n1[feat]
            blocks.append((t.float(), 1 if t.dim() < 2 else int(t.shape[-1])))
        for feat in ("charge_model", "atomic_number"):
            if feat in self.in_feat_name:
                dim = self.in_feat_dims[feat]
                t = n1[feat].float()
                if t.shape[-1] == dim:
                    blocks.append((t, dim))
                elif t.shape[-1] < dim:
                    blocks.append((None, dim - t.shape[-1]))     # reference quirk (:168-169): zeros only
                else:
                    blocks.append((t[:, :dim], dim))
        if len(blocks) > 8:
            _unsupported("more than 8 input feature blocks")
        return blocks

    def forward(self, g, plan=None):
        """graph_attention.py:143-196."""
        n1 = g.nodes["n1"].data
        plan = plan if plan is not None else get_plan(g)
        x = ops.featurize(self._feature_blocks(n1), n1["partial_charge"].float())
        if x.shape[1] != self.in_feats:
            raise AssertionError(f"the input features have width {x.shape[1]}, the model expects {self.in_feats}")
        lin = self.pre_dense[0]
        p0 = self.initial_dropout.p if self.training else 0.0
        h = ops.linear(x, lin.weight, lin.bias, act=ops.ACT_ELU, drop_p=p0)
        # training driver: hook(i, 0) before block i reads its attention parameters, hook(i, 1) before its self-interaction
        hook = getattr(self, "stage_hook", None)
        if not self.no_convs:
            for i, block in enumerate(self.blocks):
                if hook is not None:
                    hook(i, 0)
                    h = block(plan, h, pre_ffn=lambda i=i: hook(i, 1))
                else:
                    h = block(plan, h)
        lin = self.post_dense[0]
        pf = self.final_dropout.p if self.training else 0.0
        h = ops.linear(h, lin.weight, lin.bias, drop_p=pf)
        g.nodes["n1"].data["h"] = h
        return g


# --------------------------------------------------------------------------------------------------- readout ----

class FeedForwardLayer(nn.Module):
    def __init__(self, in_feats, hidden_feats=None, out_feats=None, dropout=0., skip=False, layer_norm=True):
        super().__init__()
        hidden_feats = in_feats if hidden_feats is None else hidden_feats
        out_feats = in_feats if out_feats is None else out_feats
        if not layer_norm:
            _unsupported("layer_norm=False")
        self.in_feats, self.hidden_feats, self.out_feats = in_feats, hidden_feats, out_feats
        self.linear1 = nn.Linear(in_feats, hidden_feats)
        self.linear2 = nn.Linear(hidden_feats, out_feats)
        self.dropout = nn.Dropout(dropout)
        self.activation = nn.ELU()
        self.skip = skip
        self.layer_norm = layer_norm
        self.norm1 = nn.LayerNorm(in_feats)
        assert (out_feats / in_feats).is_integer() or not skip
        if skip and out_feats != in_feats:
            _unsupported("feed-forward skip with out_feats != in_feats")

    def forward(self, x, extra_residual=None):
        """network_utils.py:44-54.  `extra_residual` is added in the same GEMM epilogue (callers' outer skips)."""
        p = self.dropout.p if self.training else 0.0
        n = ops.layer_norm(x, self.norm1.weight, self.norm1.bias)
        return ops.elu_mlp(n, self.linear1.weight, self.linear1.bias, self.linear2.weight, self.linear2.bias,
                           res1=n if self.skip else None, res2=extra_residual, drop_p=p)


class _MHAParams(nn.Module):
    """Parameter holder with nn.MultiheadAttention's names and initialisation (in_proj_weight/bias, out_proj)."""
    def __init__(self, embed_dim, num_heads):
        super().__init__()
        self.embed_dim, self.num_heads = embed_dim, num_heads
        self.in_proj_weight = nn.Parameter(torch.empty((3 * embed_dim, embed_dim)))
        self.in_proj_bias = nn.Parameter(torch.zeros(3 * embed_dim))
        self.out_proj = nn.Linear(embed_dim, embed_dim)
        nn.init.xavier_uniform_(self.in_proj_weight)
        nn.init.constant_(self.out_proj.bias, 0.)


class DottedAttWithMLP(nn.Module):
    def __init__(self, n_feats, num_heads, hidden_feats=None, layer_norm=True, dropout=0.):
        super().__init__()
        hidden_feats = n_feats * 4 if hidden_feats is None else hidden_feats
        if not layer_norm:
            _unsupported("layer_norm=False")
        if n_feats % num_heads != 0 or n_feats // num_heads != 64:
            _unsupported(f"tuple attention with head_dim {n_feats / num_heads} (kernels are specialised for 64)")
        self.layer_norm = layer_norm
        self.norm1 = nn.LayerNorm(n_feats)
        self.norm2 = nn.LayerNorm(n_feats)
        self.attn = _MHAParams(n_feats, num_heads)
        self.dropout = nn.Dropout(dropout)
        self.ff = FeedForwardLayer(n_feats, hidden_feats, out_feats=n_feats, dropout=dropout, skip=True, layer_norm=layer_norm)
        self.n_feats, self.num_heads = n_feats, num_heads

    def forward(self, x, T, n_seq, outer_residual=None, outer_feats=0):
        """network_utils.py:113-136 on tuple-major tokens (T*n_seq, n_feats).
        outer_residual: the transformer input X; when given, X * (col < outer_feats ? 2 : 1) is added to the block
        output in the last GEMM epilogue (perm_equiv_transformer.py:165 and :75-77)."""
        x, qkv = self.pre(x)
        return self.post(x, qkv, T, n_seq, outer_residual, outer_feats)

    def pre(self, x):
        """norm1 + the QKV projection (network_utils.py:119-123): row-wise, so callers may apply it to the table of
        distinct token contents instead of the tokens."""
        x = ops.layer_norm(x, self.norm1.weight, self.norm1.bias)
        return x, ops.linear(x, self.attn.in_proj_weight, self.attn.in_proj_bias)

    def post(self, x, qkv, T, n_seq, outer_residual=None, outer_feats=0, qkv_gather=None):
        """qkv_gather = (plan, names): `x`, `qkv` (and `outer_residual`, when given) are TABLES of distinct token contents
        (one row per (atom, pe)), gathered inside the attention kernel, the out-projection's epilogue and the residual
        kernel: between the table and the feed-forward block no gathered token tensor is written."""
        p = self.dropout.p if self.training else 0.0
        o = ops.tuple_attention(qkv, T, n_seq, self.num_heads, gather=qkv_gather)
        x = ops.linear(o, self.attn.out_proj.weight, self.attn.out_proj.bias, res1=x, drop_p=p, res1_gather=qkv_gather)
        x = ops.layer_norm(x, self.norm2.weight, self.norm2.bias)
        res = x if outer_residual is None else ops.token_residual(x, outer_residual, outer_feats, gather=qkv_gather)
        return self.ff(x, extra_residual=res)


class GrappaTransformer(nn.Module):
    def __init__(self, n_feats, n_heads, hidden_feats, n_layers, out_feats, permutations, layer_norm=True, dropout=0.0,
                 positional_encoding=None):
        super().__init__()
        assert out_feats == n_feats
        self.n_layers = n_layers
        self.in_feats = n_feats
        self.norm = nn.LayerNorm(n_feats)        # registered but unused in the reference too (:126 vs :145-169)
        pe = None
        if isinstance(positional_encoding, torch.Tensor):
            pe = positional_encoding.float().clone()
        elif isinstance(positional_encoding, bool) and positional_encoding:
            n_seq = permutations.shape[1]
            perms = [p.tolist() for p in permutations]
            if n_seq == 3 and all(p in [[0, 1, 2], [2, 1, 0]] for p in perms):
                pe = torch.tensor([[0.], [1.], [0.]])
            elif n_seq == 4 and all(p in [[0, 1, 2, 3], [3, 2, 1, 0], [0, 2, 1, 3], [3, 1, 2, 0]] for p in perms):
                pe = torch.tensor([[0.], [1.], [1.], [0.]])
        if pe is not None:
            if pe.shape[1] != 1:
                _unsupported("positional encodings wider than one column")
            self.register_buffer("positional_encoding", pe.float())
        else:
            self.positional_encoding = None
        self.n_seq = permutations.shape[1]
        self.n_feats = n_feats + (0 if self.positional_encoding is None else self.positional_encoding.shape[1])
        if not (self.n_feats / n_heads).is_integer():
            raise ValueError("The number of input features cannot be divided by the number of heads")
        self.transformer = nn.Sequential(*[
            DottedAttWithMLP(n_feats=self.n_feats, num_heads=n_heads, hidden_feats=hidden_feats, layer_norm=layer_norm,
                             dropout=dropout) for _ in range(n_layers)])


class Symmetriser(nn.Module):
    def __init__(self, in_feats, out_feats, permutations, permutation_prefactors=None, hidden_feats=None, n_layers=1,
                 skip=True, layer_norm=True):
        super().__init__()
        assert n_layers >= 1
        hidden_feats = in_feats if hidden_feats is None else hidden_feats
        permutations = torch.as_tensor(permutations).int()
        self.n_perm, self.n_seq = permutations.shape
        assert torch.all(permutations[0] == torch.arange(self.n_seq).int()), "permutations must include the identity permutation at the zeroth entry."
        if permutation_prefactors is None:
            permutation_prefactors = torch.ones(self.n_perm, dtype=torch.float32)
        self.register_buffer("permutation_prefactors", torch.as_tensor(permutation_prefactors).float().view(self.n_perm, 1, 1))
        self.register_buffer("permutations", permutations)
        self.n_feats, self.out_feats = in_feats, out_feats
        self.mlp = nn.Sequential(
            FeedForwardLayer(in_feats=in_feats * self.n_seq, out_feats=hidden_feats if n_layers > 1 else out_feats,
                             hidden_feats=hidden_feats, skip=False, layer_norm=layer_norm, dropout=0.),
            *[FeedForwardLayer(in_feats=hidden_feats, out_feats=hidden_feats if i != n_layers - 1 else out_feats,
                               hidden_feats=hidden_feats, skip=skip if i != n_layers - 1 else False, layer_norm=layer_norm,
                               dropout=0.) for i in range(1, n_layers)])
        # host copies for the kernels (buffers are constants of the architecture)
        self._perms_host = [[int(v) for v in p] for p in permutations.tolist()]
        self._pref_host = [float(v) for v in self.permutation_prefactors.flatten().tolist()]

    def forward(self, x, T):
        """perm_equiv_transformer.py:257-294 up to (not including) the prefactor-weighted sum over permutations,
        which is fused with the range maps in the writers.  Returns (n_perm*T, out_feats)."""
        if FOLD_SYMMETRISER and not torch.is_grad_enabled() and T > 0:
            return self._forward_folded(x, T)
        first = self.mlp[0]
        if FUSE_SYMM_LN and T > 0 and not first.skip and x.shape[1] % 4 == 0:
            # permuted copies + the first layer's LayerNorm in one kernel (statistics once per tuple, no un-normalised copy)
            n = ops.symm_layer_norm(x, first.norm1.weight, first.norm1.bias, T, self.n_seq, self._perms_host)
            rows = ops.elu_mlp(n, first.linear1.weight, first.linear1.bias, first.linear2.weight, first.linear2.bias,
                               drop_p=first.dropout.p if self.training else 0.0)
            rest = list(self.mlp)[1:]
        else:
            rows = ops.symm_permute(x, T, self.n_seq, self._perms_host)
            rest = list(self.mlp)
        for layer in rest:
            rows = layer(rows)
        return rows

    # ---- inference: permutations folded into the first layer's weights -------------------------------------------------
    def _folded_weights(self):
        """All permutations of a tuple share the LayerNorm statistics of the first layer (same multiset of values), so
        LN(x_perm; gamma, beta) = gamma (.) P x_hat + beta with x_hat the normalised un-permuted tokens, and
        W1 LN(x_perm) + b1 = (W1 diag(gamma) P) x_hat + (W1 beta + b1).  One weight view per permutation, cached per
        weight version."""
        ff = self.mlp[0]
        ps = (ff.norm1.weight, ff.norm1.bias, ff.linear1.weight, ff.linear1.bias)
        stamp = tuple((q._version, q.data_ptr()) for q in ps) + (ops.WEIGHT_EPOCH[0],)
        c = getattr(self, "_fold", None)
        if c is not None and c[0] == stamp:
            return c[1]
        g, b, w1, b1 = (q.detach() for q in ps)
        H, D = w1.shape
        E = D // self.n_seq
        wv = (w1 * g[None, :]).view(H, self.n_seq, E)
        views = []
        for perm in self._perms_host:                   # x_perm block i = x block perm[i]  ->  column block j takes block i with perm[i] = j
            inv = [perm.index(j) for j in range(self.n_seq)]
            wp = wv[:, inv, :].reshape(H, D).contiguous()
            if ops.MATH_MODE == 2:
                views.append((wp, ops.make_planes(wp)))
            else:
                views.append((wp, ops.split_lo(wp) if ops.USE_WEIGHT_LO and ops.MATH_MODE == 1 else None))
        bias = torch.addmv(b1, w1, b).contiguous()
        ones, zeros = torch.ones_like(g), torch.zeros_like(g)
        out = (views, bias, ones, zeros)
        self._fold = (stamp, out)
        return out

    def _forward_folded(self, x, T):
        ff = self.mlp[0]
        views, bias, ones, zeros = self._folded_weights()
        H, D = ff.linear1.weight.shape
        xh = ops.layer_norm(x.view(T, D), ones, zeros)                      # x_hat: statistics once per tuple
        z = torch.empty((len(views) * T, H), device=x.device, dtype=torch.float32)
        for p_i, (wp, lo) in enumerate(views):                              # rows [p T, (p+1) T) = permutation p
            if isinstance(lo, ops.WeightPlanes) and T >= 32:
                ops.gemm_bf16(T, H, D, xh, D, lo.nt, z[p_i * T:(p_i + 1) * T], H, bias=bias, act=ops.ACT_ELU)
            else:
                ops.gemm(0, T, H, D, xh, D, wp, D, z[p_i * T:(p_i + 1) * T], H, bias=bias, act=ops.ACT_ELU,
                         b_lo=None if isinstance(lo, ops.WeightPlanes) else lo)
        rows = ops.linear(z, ff.linear2.weight, ff.linear2.bias)
        for layer in list(self.mlp)[1:]:
            rows = layer(rows)
        return rows


class SymmetrisedTransformer(nn.Module):
    def __init__(self, n_feats, n_heads, hidden_feats, n_layers, out_feats, permutations, layer_norm=True, dropout=0.0,
                 symmetriser_layers=1, symmetriser_hidden_feats=None, permutation_prefactors=None, positional_encoding=None):
        super().__init__()
        if n_layers < 1:
            _unsupported("symmetric_transformer_depth = 0")
        if not layer_norm:
            _unsupported("layer_norm=False")
        self.n_feats = n_feats
        self.layer_norm = layer_norm
        self.grappa_transformer = GrappaTransformer(n_feats=n_feats, n_heads=n_heads, hidden_feats=hidden_feats,
                                                    n_layers=n_layers, out_feats=n_feats, permutations=permutations,
                                                    layer_norm=layer_norm, dropout=dropout,
                                                    positional_encoding=positional_encoding)
        self.trafo_out_feats = self.grappa_transformer.n_feats
        self.norm = nn.LayerNorm(self.trafo_out_feats)
        self.symmetriser = Symmetriser(in_feats=self.trafo_out_feats, out_feats=out_feats, permutations=permutations,
                                       permutation_prefactors=permutation_prefactors, hidden_feats=symmetriser_hidden_feats,
                                       n_layers=symmetriser_layers, layer_norm=layer_norm)

    def forward(self, atom_feats, plan, level):
        """perm_equiv_transformer.py:67-85 on tuple-major tokens, starting from the projected ATOM features:
        the gather of RepProjector (interaction_parameters.py:184-189) happens here."""
        gt = self.grappa_transformer
        n_seq, T = gt.n_seq, plan.T[level]
        pe = gt.positional_encoding
        n_blocks = len(gt.transformer)
        names = self._dedup_tables(plan, level, atom_feats.shape[0], T * n_seq)
        if names is not None:
            # distinct token contents: the projected atom rows themselves (no PE column) or the (atom, pe) table
            table = atom_feats if pe is None else ops.pe_table(atom_feats, self.n_feats, self.trafo_out_feats)
            x1_tab, qkv_tab = gt.transformer[0].pre(table)
            rows = T * n_seq
            # with one transformer block (the default) the input residual is read straight off the table inside the
            # residual kernel; deeper stacks need the gathered input again after the last block
            x_in = table if n_blocks == 1 else ops.row_gather(table, plan, names, rows)
            x = gt.transformer[0].post(x1_tab, qkv_tab, T, n_seq,
                                       outer_residual=x_in if n_blocks == 1 else None, outer_feats=self.n_feats,
                                       qkv_gather=(plan, names))
            first = 1
        else:
            x_in = ops.tuple_gather(atom_feats, plan, level, n_seq, self.n_feats, self.trafo_out_feats,
                                    None if pe is None else pe.reshape(-1).contiguous())
            x = x_in
            first = 0
        for i in range(first, n_blocks):
            last = i == n_blocks - 1
            x = gt.transformer[i](x, T, n_seq, outer_residual=x_in if last else None, outer_feats=self.n_feats)
        x = ops.layer_norm(x, self.norm.weight, self.norm.bias)
        return self.symmetriser(x, T)


def _dedup_tables(self, plan, level, n_atoms, n_tokens):
    """Plan array names (idx, ptr, tok) of the distinct-content table of `level`, or None when de-duplication does not pay
    or the positional encoding is not the pattern the plan was built for."""
    if not DEDUP_TOKENS or n_tokens == 0:
        return None
    pe = self.grappa_transformer.positional_encoding
    if pe is None:
        if self.trafo_out_feats != self.n_feats or n_tokens < DEDUP_RATIO * n_atoms:
            return None
        return ("idx_" + level, "inc_ptr_" + level, "inc_tok_" + level)
    pattern = PE_PATTERN.get(level)
    if pattern is None or ("tab_idx_" + level) not in plan.offsets or n_tokens < DEDUP_RATIO * 2 * n_atoms:
        return None
    if getattr(self, "_pe_host", None) is None:
        self._pe_host = tuple(float(v) for v in pe.detach().flatten().tolist())
    if self._pe_host != tuple(float(v) for v in pattern):
        return None
    return ("tab_idx_" + level, "tab_ptr_" + level, "tab_tok_" + level)


SymmetrisedTransformer._dedup_tables = _dedup_tables


class ToPositive(nn.Module):
    """final_layer.py:11-52; evaluated inside the readout-finalize kernel."""
    def __init__(self, mean, std, min_=0., learnable_statistics=False):
        super().__init__()
        if learnable_statistics:
            _unsupported("learnable_statistics=True")
        self.register_buffer("mean_over_std", torch.tensor(float(mean / std)))
        self.register_buffer("std", torch.tensor(float(std)))
        self.register_buffer("min_", torch.tensor(float(min_)))


class ToRange(nn.Module):
    """final_layer.py:54-97; evaluated inside the readout-finalize kernel."""
    def __init__(self, max_, std, learnable_statistics=False):
        super().__init__()
        if learnable_statistics:
            _unsupported("learnable_statistics=True")
        self.register_buffer("std_over_max", torch.tensor(float(std / max_)).float())
        self.register_buffer("max", torch.tensor(float(max_)).float())


class RepProjector(nn.Module):
    def __init__(self, dim_tupel, in_feats, out_feats, improper=False):
        super().__init__()
        self.dim_tupel, self.improper, self.out_feats = dim_tupel, improper, out_feats
        self.mlp = nn.Sequential(nn.Linear(in_feats, out_feats), nn.ELU())

    def forward(self, h):
        """interaction_parameters.py:171-172 (per-atom part only; rows padded to a multiple of 4 columns)."""
        lin = self.mlp[0]
        return ops.linear(h, lin.weight, lin.bias, act=ops.ACT_ELU, out_ld=(self.out_feats + 3) // 4 * 4)


class _HostConsts:
    """Scalar buffers are architecture constants; they are mirrored on the host once (and again after
    load_state_dict / .to()) so that launching the finalize kernel never synchronises the device."""
    def _refresh_consts(self):
        raise NotImplementedError

    def _apply(self, fn, *a, **k):
        out = super()._apply(fn, *a, **k)
        self._consts = None
        return out

    def _load_from_state_dict(self, *a, **k):
        super()._load_from_state_dict(*a, **k)
        self._consts = None

    def consts(self):
        if getattr(self, "_consts", None) is None:
            self._consts = self._refresh_consts()
        return self._consts


class WriteBondParameters(_HostConsts, nn.Module):
    LEVEL, KIND = "n2", 0

    def __init__(self, rep_feats, between_feats, suffix="", param_statistics=None, n_att=2, n_heads=8, dense_layers=2,
                 dropout=0., layer_norm=True, symmetriser_feats=None, attention_hidden_feats=None, positional_encoding=True,
                 learnable_statistics=False, gate=False, shifted_elu=True, stat_scaling=True):
        super().__init__()
        if gate:
            _unsupported("harmonic_gate=True")
        if not shifted_elu:
            _unsupported("shifted_elu=False")
        eps = 1e-6
        st = param_statistics if param_statistics is not None else get_default_statistics()
        if stat_scaling:
            k_mean, k_std = st["mean"]["n2_k"].item(), st["std"]["n2_k"].item() + eps
            eq_mean, eq_std = st["mean"]["n2_eq"].item(), st["std"]["n2_eq"].item() + eps
        else:
            k_mean, k_std, eq_mean, eq_std = 0., 1., 0., 1.
        self.suffix = suffix
        self.rep_projector = RepProjector(dim_tupel=2, in_feats=rep_feats, out_feats=between_feats)
        symmetriser_feats = between_feats if symmetriser_feats is None else symmetriser_feats
        attention_hidden_feats = 4 * between_feats if attention_hidden_feats is None else attention_hidden_feats
        self.gate = gate
        self.bond_model = SymmetrisedTransformer(
            n_feats=between_feats, n_heads=n_heads, hidden_feats=attention_hidden_feats, n_layers=n_att, out_feats=2,
            permutations=torch.tensor([[0, 1], [1, 0]], dtype=torch.int32), layer_norm=layer_norm, dropout=dropout,
            symmetriser_layers=dense_layers, symmetriser_hidden_feats=symmetriser_feats, positional_encoding=False)
        self.to_k = ToPositive(mean=k_mean, std=k_std, min_=0, learnable_statistics=learnable_statistics)
        self.to_eq = ToPositive(mean=eq_mean, std=eq_std, learnable_statistics=learnable_statistics)

    def _refresh_consts(self):
        return [self.to_eq.mean_over_std.item(), self.to_eq.std.item(), self.to_eq.min_.item(),
                self.to_k.mean_over_std.item(), self.to_k.std.item(), self.to_k.min_.item(), 0.0]

    def forward(self, g, plan):
        """interaction_parameters.py:268-290."""
        T = plan.T["n2"]
        h = g.nodes["n1"].data["h"]
        if T == 0:
            z = torch.zeros((0,), device=h.device)
            g.nodes["n2"].data["eq" + self.suffix], g.nodes["n2"].data["k" + self.suffix] = z, z.clone()
            return g
        a = self.rep_projector(h)
        s = self.bond_model(a, plan, "n2")
        eq, k = ops.readout_finalize(s, 0, T, self.bond_model.symmetriser._pref_host, self.consts())
        g.nodes["n2"].data["eq" + self.suffix] = eq
        g.nodes["n2"].data["k" + self.suffix] = k
        return g


class WriteAngleParameters(_HostConsts, nn.Module):
    def __init__(self, rep_feats, between_feats, suffix="", param_statistics=None, n_att=2, n_heads=8, dense_layers=2,
                 dropout=0., layer_norm=True, symmetriser_feats=None, attention_hidden_feats=None, positional_encoding=True,
                 learnable_statistics=False, gate=False, shifted_elu=True, stat_scaling=True):
        super().__init__()
        if gate:
            _unsupported("harmonic_gate=True")
        if not shifted_elu:
            _unsupported("shifted_elu=False")
        eps = 1e-6
        if stat_scaling:
            st = param_statistics if param_statistics is not None else get_default_statistics()
            k_mean, k_std = st["mean"]["n3_k"].item(), st["std"]["n3_k"].item() + eps
            eq_std = st["std"]["n3_eq"].item() + eps
        else:
            k_mean, k_std, eq_std = 0., 1., 1.
        self.suffix = suffix
        rep_projected_feats = between_feats if not positional_encoding else between_feats - 1
        self.rep_projector = RepProjector(dim_tupel=3, in_feats=rep_feats, out_feats=rep_projected_feats)
        symmetriser_feats = between_feats if symmetriser_feats is None else symmetriser_feats
        attention_hidden_feats = 4 * between_feats if attention_hidden_feats is None else attention_hidden_feats
        self.gate = gate
        self.angle_model = SymmetrisedTransformer(
            n_feats=rep_projected_feats, n_heads=n_heads, hidden_feats=attention_hidden_feats, n_layers=n_att, out_feats=2,
            permutations=torch.tensor([[0, 1, 2], [2, 1, 0]], dtype=torch.int32), layer_norm=layer_norm, dropout=dropout,
            symmetriser_layers=dense_layers, symmetriser_hidden_feats=symmetriser_feats,
            positional_encoding=copy.deepcopy(positional_encoding))
        self.to_k = ToPositive(mean=k_mean, std=k_std, min_=0, learnable_statistics=learnable_statistics)
        self.to_eq = ToRange(max_=math.pi, std=eq_std, learnable_statistics=learnable_statistics)

    def _refresh_consts(self):
        return [self.to_eq.std_over_max.item(), self.to_eq.max.item(), 0.0,
                self.to_k.mean_over_std.item(), self.to_k.std.item(), self.to_k.min_.item(), 0.0]

    def forward(self, g, plan):
        """interaction_parameters.py:372-397."""
        if "n3" not in g.ntypes:
            return g
        T = plan.T["n3"]
        h = g.nodes["n1"].data["h"]
        if T == 0:
            z = torch.zeros((0,), device=h.device)
            g.nodes["n3"].data["eq" + self.suffix], g.nodes["n3"].data["k" + self.suffix] = z, z.clone()
            return g
        a = self.rep_projector(h)
        s = self.angle_model(a, plan, "n3")
        eq, k = ops.readout_finalize(s, 1, T, self.angle_model.symmetriser._pref_host, self.consts())
        g.nodes["n3"].data["eq" + self.suffix] = eq
        g.nodes["n3"].data["k" + self.suffix] = k
        return g


class WriteTorsionParameters(_HostConsts, nn.Module):
    def __init__(self, rep_feats, between_feats, suffix="", n_periodicity=None, improper=False, n_att=2, n_heads=8,
                 dense_layers=2, dropout=0., layer_norm=True, symmetriser_feats=None, attention_hidden_feats=None,
                 param_statistics=None, positional_encoding=True, gated=False, learnable_statistics=False,
                 wrong_symmetry=False, cutoff=1e-4, only_n2=False, stat_scaling=True):
        super().__init__()
        if learnable_statistics:
            _unsupported("learnable_statistics=True")
        if wrong_symmetry:
            _unsupported("wrong_symmetry=True (ablation)")
        eps = 1e-1 if gated else 1e-2
        self.gated = gated
        if n_periodicity is None:
            n_periodicity = 6
        self.register_buffer("n_periodicity", torch.tensor(n_periodicity).long())
        if stat_scaling:
            st = param_statistics if param_statistics is not None else get_default_statistics()
            if not improper:
                k_mean, k_std = st["mean"]["n4_k"], st["std"]["n4_k"] + eps
            elif "n4_improper_k" not in st["mean"]:
                k_mean, k_std = torch.zeros(n_periodicity), torch.ones(n_periodicity)
            else:
                k_mean, k_std = st["mean"]["n4_improper_k"], st["std"]["n4_improper_k"] + eps
                if len(k_mean) < n_periodicity or len(k_std) < n_periodicity:
                    raise ValueError(f"n_periodicity is {n_periodicity} but the param_statistics contains fewer values for the improper torsion parameters.")
        else:
            k_mean, k_std = torch.zeros(n_periodicity), torch.ones(n_periodicity)
        k_mean = k_mean[:n_periodicity] if not only_n2 else k_mean[1]
        k_std = k_std[:n_periodicity] if not only_n2 else k_std[1]
        self.register_buffer("k_mean", k_mean.unsqueeze(dim=0).float().clone())
        self.register_buffer("k_std", k_std.unsqueeze(dim=0).float().clone())
        self.suffix, self.wrong_symmetry, self.improper = suffix, wrong_symmetry, improper
        rep_projected_feats = between_feats if not positional_encoding else between_feats - 1
        self.rep_projector = RepProjector(dim_tupel=4, in_feats=rep_feats, improper=improper, out_feats=rep_projected_feats)
        symmetriser_feats = between_feats if symmetriser_feats is None else symmetriser_feats
        attention_hidden_feats = 4 * between_feats if attention_hidden_feats is None else attention_hidden_feats
        perms = torch.tensor([[0, 1, 2, 3], [3, 2, 1, 0]] if not improper else [[0, 1, 2, 3], [3, 1, 2, 0]], dtype=torch.int32)
        n_out = n_periodicity if not gated else 2 * n_periodicity
        if only_n2:
            n_out = 1 if not gated else 2
            if n_periodicity < 2:
                raise ValueError("only_n2 needs n_periodicity >= 2")
        self.n_out_feats, self.only_n2 = n_out, only_n2
        self.torsion_model = SymmetrisedTransformer(
            n_feats=rep_projected_feats, n_heads=n_heads, hidden_feats=attention_hidden_feats, n_layers=n_att,
            out_feats=n_out, permutations=perms, layer_norm=layer_norm, dropout=dropout, symmetriser_layers=dense_layers,
            symmetriser_hidden_feats=symmetriser_feats, positional_encoding=copy.deepcopy(positional_encoding))
        self.cutoff_value = float(cutoff) if cutoff > 0 else 0.0

    def _refresh_consts(self):
        return [0.0, 0.0, 0.0, 0.0, 0.0, 0.0, self.cutoff_value], int(self.n_periodicity.item())

    def forward(self, g, plan):
        """interaction_parameters.py:567-615."""
        level = "n4_improper" if self.improper else "n4"
        if level not in g.ntypes:
            return g
        consts, n_per = self.consts()
        T = plan.T[level]
        h = g.nodes["n1"].data["h"]
        if T == 0:
            g.nodes[level].data["k" + self.suffix] = torch.zeros((0, n_per), device=h.device)
            return g
        a = self.rep_projector(h)
        s = self.torsion_model(a, plan, level)
        k = ops.readout_finalize(s, 2 if self.gated else 3, T, self.torsion_model.symmetriser._pref_host, consts,
                                 kstd=self.k_std.reshape(-1), kmean=self.k_mean.reshape(-1), n_per=n_per,
                                 col0=1 if self.only_n2 else 0)
        g.nodes[level].data["k" + self.suffix] = k
        return g


class WriteParameters(nn.Module):
    def __init__(self, graph_node_features=256, symmetric_transformer_dropout=0, layer_norm=True, positional_encoding=True,
                 transformer_depth=2, n_heads=8, transformer_width=512, symmetriser_depth=2, symmetriser_width=256,
                 n_periodicity_proper=6, n_periodicity_improper=3, gated_torsion=False, suffix="", wrong_symmetry=False,
                 learnable_statistics=False, param_statistics=None, torsion_cutoff=1.e-4, harmonic_gate=False,
                 only_n2_improper=False, shifted_elu=True, stat_scaling=True):
        super().__init__()
        st = param_statistics if param_statistics is not None else get_default_statistics()
        for m in ("mean", "std"):                           # interaction_parameters.py:45-49
            for k, v in st[m].items():
                if torch.isnan(torch.as_tensor(v)).any():
                    st[m][k] = get_default_statistics()[m][k]
        common = dict(rep_feats=graph_node_features, between_feats=transformer_width, suffix=suffix, n_att=transformer_depth,
                      n_heads=n_heads, dense_layers=symmetriser_depth, dropout=symmetric_transformer_dropout,
                      layer_norm=layer_norm, symmetriser_feats=symmetriser_width, attention_hidden_feats=transformer_width,
                      learnable_statistics=learnable_statistics, stat_scaling=stat_scaling, param_statistics=st)
        self.bond_writer = WriteBondParameters(gate=harmonic_gate, shifted_elu=shifted_elu, **common)
        self.angle_writer = WriteAngleParameters(positional_encoding=positional_encoding, gate=harmonic_gate,
                                                 shifted_elu=shifted_elu, **common)
        self.proper_writer = WriteTorsionParameters(n_periodicity=n_periodicity_proper, improper=False,
                                                    positional_encoding=positional_encoding, gated=gated_torsion,
                                                    cutoff=torsion_cutoff, **common)
        self.improper_writer = WriteTorsionParameters(n_periodicity=n_periodicity_improper, improper=True,
                                                      positional_encoding=positional_encoding, gated=gated_torsion,
                                                      wrong_symmetry=wrong_symmetry, cutoff=torsion_cutoff,
                                                      only_n2=only_n2_improper, **common)

    def forward(self, g, plan):
        """interaction_parameters.py:136-146.  The four writers only share their input (the atom embedding), so they run
        on four parallel streams (forked from / joined into the caller's stream; under CUDA-graph capture these become
        parallel branches).  Their GEMMs have 20-100 tiles each, far fewer than the 148 SMs, so running them side by
        side is what fills the machine.  Autograd replays each writer's backward on the stream of its forward."""
        writers = (self.bond_writer, self.angle_writer, self.proper_writer, self.improper_writer)
        h = g.nodes["n1"].data["h"]
        if not (PARALLEL_WRITERS and h.is_cuda):
            for w in writers:
                g = w(g, plan)
            return g
        main = torch.cuda.current_stream(h.device)
        streams = _writer_streams(h.device)
        fork = torch.cuda.Event()
        fork.record(main)
        for w, st in zip(writers, streams):
            st.wait_event(fork)
            with torch.cuda.stream(st):
                g = w(g, plan)
        for st in streams:
            main.wait_stream(st)
        return g


# ------------------------------------------------------------------------------------------------ GrappaModel ---

class GrappaModel(nn.Module):
    def __init__(self, graph_node_features: int = 256, in_feats: int = None,
                 in_feat_name: Union[str, List[str]] = ["atomic_number", "ring_encoding", "partial_charge", "degree"],
                 in_feat_dims: Dict[str, int] = {}, gnn_width: int = 512, gnn_attentional_layers: int = 4,
                 gnn_convolutions: int = 0, gnn_attention_heads: int = 16, gnn_dropout_attention: float = 0.3,
                 gnn_dropout_initial: float = 0., gnn_dropout_conv: float = 0., gnn_dropout_final: float = 0.1,
                 symmetric_transformer_dropout: float = 0.5, symmetric_transformer_depth: int = 1,
                 symmetric_transformer_n_heads: int = 8, symmetric_transformer_width: int = 512,
                 symmetriser_depth: int = 4, symmetriser_width: int = 256, n_periodicity_proper: int = 3,
                 n_periodicity_improper: int = 3, gated_torsion: bool = False, positional_encoding: bool = True,
                 layer_norm: bool = True, self_interaction: bool = True, learnable_statistics: bool = False,
                 param_statistics: dict = None, torsion_cutoff: float = 1.e-4, harmonic_gate: bool = False,
                 only_n2_improper: bool = True, stat_scaling: bool = True, shifted_elu: bool = True):
        """Keyword-compatible with models/grappa.py:8-39 (Python defaults included: gated_torsion False and
        n_periodicity_improper 3 differ from configs/model/default.yaml; use `from_default_config()` for the YAML)."""
        super().__init__()
        if param_statistics is None:
            param_statistics = get_default_statistics()
        self.gnn = GrappaGNN(out_feats=graph_node_features, in_feats=in_feats, node_feats=gnn_width, n_conv=gnn_convolutions,
                             n_att=gnn_attentional_layers, n_heads=gnn_attention_heads, in_feat_name=in_feat_name,
                             in_feat_dims=in_feat_dims, conv_dropout=gnn_dropout_conv, attention_dropout=gnn_dropout_attention,
                             final_dropout=gnn_dropout_final, initial_dropout=gnn_dropout_initial,
                             self_interaction=self_interaction, layer_norm=layer_norm)
        self.parameter_writer = WriteParameters(
            graph_node_features=graph_node_features, symmetric_transformer_dropout=symmetric_transformer_dropout,
            layer_norm=layer_norm, positional_encoding=positional_encoding, transformer_depth=symmetric_transformer_depth,
            n_heads=symmetric_transformer_n_heads, transformer_width=symmetric_transformer_width,
            symmetriser_depth=symmetriser_depth, symmetriser_width=symmetriser_width,
            n_periodicity_proper=n_periodicity_proper, n_periodicity_improper=n_periodicity_improper, wrong_symmetry=False,
            param_statistics=param_statistics, gated_torsion=gated_torsion, learnable_statistics=learnable_statistics,
            torsion_cutoff=torsion_cutoff, harmonic_gate=harmonic_gate, only_n2_improper=only_n2_improper,
            stat_scaling=stat_scaling, shifted_elu=shifted_elu)
        self.field_of_view = gnn_attentional_layers + gnn_convolutions + 3

    @classmethod
    def from_default_config(cls, **overrides) -> "GrappaModel":
        cfg = copy.deepcopy(DEFAULT_CONFIG)
        cfg.update(overrides)
        return cls(**cfg)

    def forward(self, g):
        """models/grappa.py:168-189.  The idxs bounds check (:179-185) is done on the host when the batch plan is
        built, so it costs no device synchronisation."""
        plan = get_plan(g)
        g = self.gnn(g, plan)
        g = self.parameter_writer(g, plan)
        return g
