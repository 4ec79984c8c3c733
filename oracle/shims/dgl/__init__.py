"""Minimal DGL stand-in: homogeneous graph + the message-passing primitives used
by generic/eegatconv.py (u_add_v, u_mul_e, sum, edge_softmax)."""
import contextlib
import torch
from . import function, base, utils  # noqa: F401
from . import nn  # noqa: F401


class _EdgeBatch:
    def __init__(self, g):
        self.data = g.edata
        self.src = {k: v[g._src] for k, v in g.srcdata.items()}
        self.dst = {k: v[g._dst] for k, v in g.dstdata.items()}


class DGLGraph:
    def __init__(self, data, num_nodes=None):
        src, dst = data
        self._src = torch.as_tensor(src).long()
        self._dst = torch.as_tensor(dst).long()
        if num_nodes is None:
            num_nodes = int(max(self._src.max(), self._dst.max())) + 1
        self._n = num_nodes
        self.ndata = {}
        self.edata = {}

    # homogeneous graph: src and dst frames are the node frame
    @property
    def srcdata(self):
        return self.ndata

    @property
    def dstdata(self):
        return self.ndata

    def num_nodes(self):
        return self._n

    def num_edges(self):
        return self._src.numel()

    def edges(self):
        return self._src, self._dst

    def in_degrees(self):
        return torch.bincount(self._dst, minlength=self._n)

    @contextlib.contextmanager
    def local_scope(self):
        nd, ed = dict(self.ndata), dict(self.edata)
        try:
            yield
        finally:
            self.ndata.clear(); self.ndata.update(nd)
            self.edata.clear(); self.edata.update(ed)

    def apply_edges(self, func):
        if isinstance(func, function.BinaryMessage):
            assert func.kind == "u_op_v"
            self.edata[func.out] = func.op(self.ndata[func.lhs][self._src],
                                           self.ndata[func.rhs][self._dst])
        else:
            self.edata.update(func(_EdgeBatch(self)))

    def update_all(self, message_func, reduce_func):
        assert message_func.kind == "u_op_e"
        m = message_func.op(self.ndata[message_func.lhs][self._src],
                            self.edata[message_func.rhs])
        assert reduce_func.name == "sum"
        out = torch.zeros((self._n,) + tuple(m.shape[1:]), dtype=m.dtype)
        self.ndata[reduce_func.out] = out.index_add(0, self._dst, m)


def graph(data, num_nodes=None):
    return DGLGraph(data, num_nodes)
