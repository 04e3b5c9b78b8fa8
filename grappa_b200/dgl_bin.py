"""
DGL-free reader for heterograph files written by `dgl.save_graphs`, enough to load Grappa's
datasets (`data/dataset.py:171-222` reads them with `dgl.load_graphs`) without DGL: node data of every node type and the
type names.  The bond edges are NOT read from the file's graph structure: Grappa's graphs carry their bonds as the
`idxs` of the `n2` nodes and build the `n1` edges from them (`data/molecule.py:530-537`), which is what
`to_heterograph` does.

File layout (DGL 2.x `graph_serialize` version 2, heterograph flavour, dmlc little-endian serialisation), as far as
this reader needs it:
    u64 magic 0xDD2E4FF046B4A13F | u64 version = 2 | u64 graph type = 2 (hetero) | u64 number of graphs | padding to 4096
    ... index / label section ...
    per graph, starting at the heterograph magic 0xDD589FBE35224ABF:
        graph structure (meta graph + one unit graph per relation)              -- skipped
        node tensors   vector<vector<pair<string, NDArray>>>, one inner vector per node type (types sorted by name)
        edge tensors   same shape, per edge type
        node type names vector<string>, edge type names vector<string>
    NDArray: u64 magic 0xDD5E40F096B4A13F | u64 reserved | i32 device type | i32 device id | i32 ndim |
             u8 dtype code (0 int, 1 uint, 2 float) | u8 bits | u16 lanes | i64 shape[ndim] | i64 nbytes | data
The node-tensor section is located by scanning for the first length-prefixed feature name inside the graph's block and
checked by parsing through to the type-name lists, which must end exactly where the next graph starts.
"""
import re
import struct
from typing import Dict, List, Tuple

import numpy as np

FILE_MAGIC = 0xDD2E4FF046B4A13F
HETERO_MAGIC = 0xDD589FBE35224ABF
NDARRAY_MAGIC = 0xDD5E40F096B4A13F

_DTYPES = {(0, 8): np.int8, (0, 16): np.int16, (0, 32): np.int32, (0, 64): np.int64,
           (1, 8): np.uint8, (1, 16): np.uint16, (1, 32): np.uint32, (1, 64): np.uint64,
           (2, 16): np.float16, (2, 32): np.float32, (2, 64): np.float64}


class _Reader:
    def __init__(self, buf: bytes, pos: int):
        self.b, self.p = buf, pos

    def u64(self) -> int:
        v = struct.unpack_from("<Q", self.b, self.p)[0]
        self.p += 8
        return v

    def string(self) -> str:
        n = self.u64()
        if n > 4096:
            raise ValueError(f"implausible string length {n} at offset {self.p - 8}")
        s = self.b[self.p:self.p + n].decode("ascii")
        self.p += n
        return s

    def ndarray(self) -> np.ndarray:
        magic = self.u64()
        if magic != NDARRAY_MAGIC:
            raise ValueError(f"expected an NDArray at offset {self.p - 8}")
        self.p += 8                                               # reserved
        _dev_type, _dev_id, ndim = struct.unpack_from("<iii", self.b, self.p)
        self.p += 12
        code, bits, lanes = struct.unpack_from("<BBH", self.b, self.p)
        self.p += 4
        if lanes != 1 or (code, bits) not in _DTYPES or not 0 <= ndim <= 8:
            raise ValueError(f"unsupported NDArray header (code {code}, bits {bits}, lanes {lanes}, ndim {ndim})")
        shape = struct.unpack_from(f"<{ndim}q", self.b, self.p)
        self.p += 8 * ndim
        nbytes = struct.unpack_from("<q", self.b, self.p)[0]
        self.p += 8
        dt = np.dtype(_DTYPES[(code, bits)])
        count = int(np.prod(shape)) if ndim else 1
        if nbytes != count * dt.itemsize:
            raise ValueError("NDArray byte size does not match its shape")
        a = np.frombuffer(self.b, dtype=dt, count=count, offset=self.p).reshape(shape).copy()
        self.p += nbytes
        return a

    def tensor_lists(self) -> List[Dict[str, np.ndarray]]:
        out = []
        for _ in range(self.u64()):
            d = {}
            for _ in range(self.u64()):
                name = self.string()
                d[name] = self.ndarray()
            out.append(d)
        return out

    def strings(self) -> List[str]:
        return [self.string() for _ in range(self.u64())]


def load_graphs(path: str) -> List[Dict[str, Dict[str, np.ndarray]]]:
    """Returns one {node type: {feature name: array}} per graph in the file, in file order."""
    with open(path, "rb") as f:
        buf = f.read()
    magic, version, gtype, n_graphs = struct.unpack_from("<4Q", buf, 0)
    if magic != FILE_MAGIC:
        raise ValueError("not a DGL graph file (bad magic)")
    if version != 2 or gtype != 2:
        raise ValueError(f"unsupported DGL graph file (version {version}, graph type {gtype}); only heterograph files of version 2")
    starts = [m.start() for m in re.finditer(re.escape(struct.pack("<Q", HETERO_MAGIC)), buf)]
    starts = [s for s in starts if s >= 4096]
    if len(starts) != n_graphs:
        raise ValueError(f"found {len(starts)} graph records, header says {n_graphs}")
    ends = starts[1:] + [len(buf)]
    name_pat = re.compile(rb"([\x01-\x40])\x00{7}([A-Za-z_][A-Za-z0-9_]*)")
    graphs = []
    for s, e in zip(starts, ends):
        parsed = None
        for m in name_pat.finditer(buf, s, e):
            if len(m.group(2)) < m.group(1)[0]:
                continue
            start = m.start() - 16                                # outer vector size, inner vector size, then the name
            if start < s:
                continue
            r = _Reader(buf, start)
            try:
                node_tensors = r.tensor_lists()
                edge_tensors = r.tensor_lists()
                ntypes, etypes = r.strings(), r.strings()
            except (ValueError, struct.error, UnicodeDecodeError):
                continue
            # records are contiguous; only the last one may be followed by a trailer (the file's label section)
            at_end = r.p == e or (e == len(buf) and r.p <= e)
            if at_end and len(ntypes) == len(node_tensors) and len(etypes) == len(edge_tensors):
                parsed = dict(zip(ntypes, node_tensors))
                break
        if parsed is None:
            raise ValueError(f"could not locate the node tensors of the graph record at offset {s}")
        graphs.append(parsed)
    return graphs


def to_heterograph(node_data: Dict[str, Dict[str, np.ndarray]]):
    """{node type: {feature: array}} of one molecule -> grappa_b200.graph.HeteroGraph (bond edges from the n2 idxs, ordered
    [all i->j, then all j->i] as data/molecule.py:530-537)."""
    import torch
    from .graph import HeteroGraph
    bonds = np.asarray(node_data["n2"]["idxs"], dtype=np.int64).reshape(-1, 2)
    src = np.concatenate([bonds[:, 0], bonds[:, 1]])
    dst = np.concatenate([bonds[:, 1], bonds[:, 0]])
    counts = {}
    for nt, feats in node_data.items():
        counts[nt] = int(next(iter(feats.values())).shape[0]) if feats else (1 if nt == "g" else 0)
    g = HeteroGraph(counts, torch.as_tensor(np.stack([src, dst]).astype(np.int32)))
    for nt, feats in node_data.items():
        for k, v in feats.items():
            g.nodes[nt].data[k] = torch.as_tensor(v)
    return g
