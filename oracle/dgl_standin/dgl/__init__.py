"""
TEST INFRASTRUCTURE ONLY — a pure-PyTorch stand-in for the handful of DGL 2.4.0 entry points that the
reference's hot-path files call.  It exists so that the *unmodified* reference files
under /root/reference/src can be imported in the build container (DGL is not installable there) to
generate the golden vectors in tests/golden/.  It is never imported by the product package `grappa_b200`.

Semantics restated from DGL 2.4.0's public Python sources (python/dgl/heterograph.py, batch.py,
readout.py, nn/pytorch/conv/dotgatconv.py); no DGL code is copied.  "parity unpinned" for DotGatConv:
no reference test pins its numerics.
"""
import sys
import types
import torch

__version__ = "2.4.0-standin"


class _NodeView:
    def __init__(self, g, nt):
        self._g, self._nt = g, nt

    @property
    def data(self):
        return self._g._ndata[self._nt]


class _NodesAccessor:
    def __init__(self, g):
        self._g = g

    def __getitem__(self, nt):
        if nt not in self._g._ndata:
            raise KeyError(nt)
        return _NodeView(self._g, nt)


class _NDataView:
    """g.ndata[feat] -> {ntype: tensor} (heterograph flavour) ; homogeneous: g.ndata[feat] -> tensor."""
    def __init__(self, g):
        self._g = g

    def __getitem__(self, feat):
        if len(self._g.ntypes) == 1:
            return self._g._ndata[self._g.ntypes[0]][feat]
        return {nt: d[feat] for nt, d in self._g._ndata.items() if feat in d}

    def __setitem__(self, feat, val):
        if len(self._g.ntypes) == 1:
            self._g._ndata[self._g.ntypes[0]][feat] = val
        else:
            for nt, v in val.items():
                self._g._ndata[nt][feat] = v

    def keys(self):
        ks = []
        for d in self._g._ndata.values():
            for k in d.keys():
                if k not in ks:
                    ks.append(k)
        return ks

    def __contains__(self, feat):
        return feat in self.keys()


class DGLGraph:
    """Minimal heterograph container: per-type node counts, per-type node data dicts, typed edge lists."""

    def __init__(self, num_nodes, edges, batch_num_nodes=None, batch_num_edges=None):
        # num_nodes: {ntype: int}; edges: {(src_t, et, dst_t): (src, dst)}
        self._num_nodes = dict(num_nodes)
        self._edges = {k: (v[0], v[1]) for k, v in edges.items()}
        self._ndata = {nt: {} for nt in self._num_nodes}
        self._batch_num_nodes = batch_num_nodes  # {ntype: LongTensor} or None
        self._batch_num_edges = batch_num_edges

    # ---- structure ----
    @property
    def ntypes(self):
        return sorted(self._num_nodes.keys())

    @property
    def etypes(self):
        return [k[1] for k in self._edges.keys()]

    @property
    def canonical_etypes(self):
        return list(self._edges.keys())

    def num_nodes(self, ntype=None):
        if ntype is None:
            return sum(self._num_nodes.values())
        return self._num_nodes[ntype]

    number_of_nodes = num_nodes

    def _ekey(self, etype):
        if etype is None:
            assert len(self._edges) == 1
            return next(iter(self._edges))
        if isinstance(etype, tuple):
            return etype
        for k in self._edges:
            if k[1] == etype:
                return k
        raise KeyError(etype)

    def num_edges(self, etype=None):
        if etype is None:
            return sum(len(v[0]) for v in self._edges.values())
        return len(self._edges[self._ekey(etype)][0])

    def edges(self, etype=None):
        return self._edges[self._ekey(etype)]

    @property
    def nodes(self):
        return _NodesAccessor(self)

    @property
    def ndata(self):
        return _NDataView(self)

    @property
    def device(self):
        for d in self._ndata.values():
            for v in d.values():
                return v.device
        for s, _ in self._edges.values():
            return s.device
        return torch.device("cpu")

    @property
    def batch_size(self):
        if self._batch_num_nodes is None:
            return 1
        return len(next(iter(self._batch_num_nodes.values())))

    def batch_num_nodes(self, ntype=None):
        if ntype is None:
            assert len(self._num_nodes) == 1
            ntype = self.ntypes[0]
        if self._batch_num_nodes is None:
            return torch.tensor([self._num_nodes[ntype]], dtype=torch.int64, device=self.device)
        return self._batch_num_nodes[ntype]

    # ---- transforms ----
    def to(self, device):
        g = DGLGraph(self._num_nodes,
                     {k: (s.to(device), d.to(device)) for k, (s, d) in self._edges.items()},
                     None if self._batch_num_nodes is None else {k: v.to(device) for k, v in self._batch_num_nodes.items()},
                     None if self._batch_num_edges is None else {k: v.to(device) for k, v in self._batch_num_edges.items()})
        for nt, d in self._ndata.items():
            for k, v in d.items():
                g._ndata[nt][k] = v.to(device)
        return g

    def node_type_subgraph(self, ntypes):
        ntypes = list(ntypes)
        g = DGLGraph({nt: self._num_nodes[nt] for nt in ntypes},
                     {k: v for k, v in self._edges.items() if k[0] in ntypes and k[2] in ntypes},
                     None if self._batch_num_nodes is None else {nt: self._batch_num_nodes[nt] for nt in ntypes},
                     None)
        for nt in ntypes:
            g._ndata[nt] = self._ndata[nt]  # shared dict, like a DGL frame view
        return g

    def to_networkx(self):
        import networkx as nx
        assert len(self._edges) == 1
        (s, d), = self._edges.values()
        nxg = nx.MultiDiGraph()
        nxg.add_nodes_from(range(self.num_nodes()))
        nxg.add_edges_from(zip(s.tolist(), d.tolist()))
        return nxg

    def subgraph(self, ids):
        assert len(self._num_nodes) == 1
        nt = self.ntypes[0]
        ids = torch.as_tensor(ids).long()
        remap = -torch.ones(self._num_nodes[nt], dtype=torch.long)
        remap[ids] = torch.arange(len(ids))
        k, (s, d) = next(iter(self._edges.items()))
        keep = (remap[s.long()] >= 0) & (remap[d.long()] >= 0)
        g = DGLGraph({nt: len(ids)}, {k: (remap[s.long()][keep], remap[d.long()][keep])})
        for f, v in self._ndata[nt].items():
            g._ndata[nt][f] = v[ids]
        return g

    def local_scope(self):
        import contextlib
        return contextlib.nullcontext()

    def __repr__(self):
        return f"DGLGraph-standin(num_nodes={self._num_nodes}, num_edges={ {k: len(v[0]) for k, v in self._edges.items()} })"


def _heterograph(data_dict, num_nodes_dict=None, idtype=None, device=None):
    num_nodes = {}
    edges = {}
    for (st, et, dt), (s, d) in data_dict.items():
        s = torch.as_tensor(s)
        d = torch.as_tensor(d)
        edges[(st, et, dt)] = (s, d)
        for t, ids in ((st, s), (dt, d)):
            n = int(ids.max().item()) + 1 if len(ids) > 0 else 0
            num_nodes[t] = max(num_nodes.get(t, 0), n)
    if num_nodes_dict is not None:
        num_nodes.update(num_nodes_dict)
    return DGLGraph(num_nodes, edges)


class _CallableModule(types.ModuleType):
    """`import dgl.heterograph` (reference: data/molecule.py:9) must leave `dgl.heterograph(...)` callable."""
    def __call__(self, *a, **kw):
        return _heterograph(*a, **kw)


heterograph = _CallableModule("dgl.heterograph")
sys.modules["dgl.heterograph"] = heterograph


def node_type_subgraph(g, ntypes):
    return g.node_type_subgraph(ntypes)


def to_homogeneous(g):
    assert len(g.ntypes) == 1, "stand-in supports single-type graphs only"
    return g


def batch(graphs):
    ntypes = graphs[0].ntypes
    num_nodes = {nt: sum(g.num_nodes(nt) for g in graphs) for nt in ntypes}
    bnn = {nt: torch.tensor([g.num_nodes(nt) for g in graphs], dtype=torch.int64) for nt in ntypes}
    edges = {}
    for k in graphs[0].canonical_etypes:
        so, do = 0, 0
        ss, ds = [], []
        for g in graphs:
            s, d = g._edges[k]
            ss.append(s + so)
            ds.append(d + do)
            so += g.num_nodes(k[0])
            do += g.num_nodes(k[2])
        edges[k] = (torch.cat(ss), torch.cat(ds))
    out = DGLGraph(num_nodes, edges, bnn)
    for nt in ntypes:
        keys = list(graphs[0]._ndata[nt].keys())
        for f in keys:
            out._ndata[nt][f] = torch.cat([g._ndata[nt][f] for g in graphs], dim=0)
    return out


def unbatch(g):
    if g._batch_num_nodes is None:
        return [g]
    n = g.batch_size
    outs = []
    offs = {nt: 0 for nt in g.ntypes}
    for i in range(n):
        nn_i = {nt: int(g._batch_num_nodes[nt][i]) for nt in g.ntypes}
        edges = {}
        for k, (s, d) in g._edges.items():
            lo_s, lo_d = offs[k[0]], offs[k[2]]
            m = (s >= lo_s) & (s < lo_s + nn_i[k[0]])
            edges[k] = (s[m] - lo_s, d[m] - lo_d)
        sg = DGLGraph(nn_i, edges)
        for nt in g.ntypes:
            for f, v in g._ndata[nt].items():
                sg._ndata[nt][f] = v[offs[nt]:offs[nt] + nn_i[nt]]
        for nt in g.ntypes:
            offs[nt] += nn_i[nt]
        outs.append(sg)
    return outs


def readout_nodes(graph, feat, weight=None, *, op="sum", ntype=None):
    assert op == "sum" and weight is None
    x = graph.nodes[ntype].data[feat]
    counts = graph.batch_num_nodes(ntype).to(x.device)
    seg = torch.repeat_interleave(torch.arange(len(counts), device=x.device), counts)
    out = torch.zeros((len(counts),) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    return out.index_add(0, seg, x)


def save_graphs(*a, **kw):
    raise NotImplementedError("dgl stand-in: save_graphs")


def load_graphs(*a, **kw):
    raise NotImplementedError("dgl stand-in: load_graphs")


from . import nn  # noqa: E402,F401
