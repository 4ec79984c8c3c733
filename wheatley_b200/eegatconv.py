"""EGATConv / EEGATConv: B200-native drop-ins for ``dgl.nn.EGATConv``
(jssp/models/gnn_dgl.py:30,166-172) and ``generic/eegatconv.py:18-313``.

Same constructor, ``forward`` signature, parameter names / shapes / init and error
behaviour; the body of ``forward`` (generic/eegatconv.py:260-292: three projections,
``u_add_v``, ``leaky_relu``, attention dot, ``edge_softmax``, ``u_mul_e``+``sum``) is one
fused CUDA kernel (``b200gat_conv_fwd``) with a fused backward."""
import torch
from torch import nn
from torch.nn import init

from . import _lib
from .dense import project3
from .edge_feats import TypedEdgeFeats
from .functional import (EdgeFeatures, attention_weights, edge_output, fused_gat_conv,
                         project_edge_feats)
from .plan import GraphPlan


class DGLError(Exception):
    """Stand-in for ``dgl.base.DGLError`` (raised on 0-in-degree nodes, eegatconv.py:238-247)."""


class LazyEdgeOutput:
    """The ``f_out [E, H, De]`` return value of EGATConv, produced on request.

    Its only caller discards it (jssp/models/gnn_dgl.py:374-379) and at 100x20-clique it
    is 210 MB per graph, so it is recomputed by ``b200gat_edge_out`` when asked for."""

    def __init__(self, fn, shape):
        self._fn, self._val, self.shape = fn, None, shape

    def materialize(self) -> torch.Tensor:
        if self._val is None:
            self._val = self._fn()
        return self._val

    def flatten(self, *a, **k):
        return self.materialize().flatten(*a, **k)

    def __torch_function__(self, func, types, args=(), kwargs=None):
        args = tuple(a.materialize() if isinstance(a, LazyEdgeOutput) else a for a in args)
        return func(*args, **(kwargs or {}))


def resolve_plan(graph):
    """Accepts a GraphPlan, an object carrying one (``.plan``), a DGL-like graph
    (``.edges()`` + ``.num_nodes()``) or a ``(src, dst, n_nodes)`` tuple."""
    if isinstance(graph, GraphPlan):
        return graph
    plan = getattr(graph, "plan", None)
    if isinstance(plan, GraphPlan):
        return plan
    cached = getattr(graph, "_b200gat_plan", None)
    if isinstance(cached, GraphPlan):
        return cached
    if isinstance(graph, tuple):
        src, dst, n = graph
    else:
        src, dst = graph.edges()
        n = graph.num_nodes()
    plan = GraphPlan(src, dst, int(n))
    try:
        graph._b200gat_plan = plan
    except AttributeError:
        pass
    return plan


class EEGATConv(nn.Module):
    def __init__(self, in_node_feats, in_edge_feats, out_node_feats, out_edge_feats, num_heads,
                 bias=True, edge_scoring=False):
        super().__init__()
        if isinstance(in_node_feats, tuple):
            raise NotImplementedError("bipartite (src,dst) feature pairs are not on Wheatley's path")
        if out_node_feats != out_edge_feats:
            raise NotImplementedError(
                "fused kernel needs out_node_feats == out_edge_feats (always true in Wheatley: "
                "jssp/models/gnn_dgl.py:166-172, psp/graph/graph_conv.py:42-49)")
        self._num_heads = num_heads
        self._in_src_node_feats = self._in_dst_node_feats = in_node_feats
        self._out_node_feats = out_node_feats
        self._out_edge_feats = out_edge_feats
        self._edge_scoring = edge_scoring
        # parameter names / shapes: generic/eegatconv.py:134-162
        self.fc_node_src = nn.Linear(in_node_feats, out_node_feats * num_heads, bias=False)
        self.fc_ni = nn.Linear(in_node_feats, out_edge_feats * num_heads, bias=False)
        self.fc_nj = nn.Linear(in_node_feats, out_edge_feats * num_heads, bias=False)
        if edge_scoring:
            self.fc_edge_src = nn.Linear(in_edge_feats, out_edge_feats * num_heads, bias=False)
        self.fc_fij = nn.Linear(in_edge_feats, out_edge_feats * num_heads, bias=False)
        self.attn = nn.Parameter(torch.empty(1, num_heads, out_edge_feats))
        if edge_scoring:
            self.fc_fij_e = nn.Linear(in_edge_feats, out_edge_feats * num_heads, bias=False)
            self.attn_e = nn.Parameter(torch.empty(1, num_heads, out_edge_feats))
        if bias:
            self.bias = nn.Parameter(torch.empty(num_heads * out_edge_feats))
        else:
            self.register_buffer("bias", None)
        self.reset_parameters()

    def reset_parameters(self):
        """generic/eegatconv.py:165-180."""
        gain = init.calculate_gain("relu")
        init.xavier_normal_(self.fc_node_src.weight, gain=gain)
        init.xavier_normal_(self.fc_ni.weight, gain=gain)
        init.xavier_normal_(self.fc_fij.weight, gain=gain)
        init.xavier_normal_(self.fc_nj.weight, gain=gain)
        init.xavier_normal_(self.attn, gain=gain)
        if self._edge_scoring:
            init.xavier_normal_(self.fc_edge_src.weight, gain=gain)
            init.xavier_normal_(self.fc_fij_e.weight, gain=gain)
            init.xavier_normal_(self.attn_e, gain=gain)
        if self.bias is not None:
            init.constant_(self.bias, 0)

    def _edge_features(self, plan, efeats, efeats_e):
        """Projected edge term (+ pre-activation bias, eegatconv.py:262,271-275)."""
        if not self._edge_scoring:
            return project_edge_feats(plan, efeats, self.fc_fij.weight, self.bias)
        if isinstance(efeats, TypedEdgeFeats):
            efeats = efeats.dense()
        dense = self.fc_fij(efeats) + self.fc_fij_e(efeats_e)
        if self.bias is not None:
            dense = dense + self.bias
        return EdgeFeatures(_lib.EDGE_DENSE, dense=dense)

    def forward(self, graph, nfeats, efeats, efeats_e=None, edge_weight=None, get_attention=False):
        if edge_weight is not None:
            raise NotImplementedError("edge_weight is never passed on Wheatley's path "
                                      "(jssp/models/gnn_dgl.py:374-376, psp/graph/graph_conv.py:55)")
        plan = resolve_plan(graph)
        if plan.stats_host()[0] > 0:                                 # eegatconv.py:238-247
            raise DGLError(
                "There are 0-in-degree nodes in the graph, output for those nodes will be invalid. "
                "Adding self-loop on the input graph by calling `g = dgl.add_self_loop(g)` will "
                "resolve the issue.")
        H, De, Do = self._num_heads, self._out_edge_feats, self._out_node_feats
        if efeats_e is None:
            efeats_e = efeats.dense() if (self._edge_scoring and isinstance(efeats, TypedEdgeFeats)) else efeats
        # :260, :261, :288 - the three bias-free projections, one tcgen05 launch when TF32 is allowed
        f_ni, f_nj, h_src = project3(nfeats, self.fc_ni.weight, self.fc_nj.weight, self.fc_node_src.weight)
        ef = self._edge_features(plan, efeats, efeats_e)
        out, saved = fused_gat_conv(plan, f_ni, f_nj, h_src, self.attn, heads=H, dim=De,
                                    flavour=_lib.FLAVOUR_ADDITIVE, slope=0.01, eps=0.0, edge_feats=ef)
        h_out = out.view(-1, H, Do)
        f_out = LazyEdgeOutput(
            lambda: edge_output(plan, f_ni, f_nj, ef, heads=H, dim=De, slope=0.01, attn=self.attn),
            (plan.n_edges, H, De))
        rets = [h_out, f_out]
        if self._edge_scoring:
            # second softmax on attn_e over the same f_out (:281-283); attention weights are
            # produced without autograd (edge scoring is dead code upstream)
            with torch.no_grad():
                _, saved_e = fused_gat_conv(plan, f_ni, f_nj, h_src, self.attn_e, heads=H, dim=De,
                                            flavour=_lib.FLAVOUR_ADDITIVE, slope=0.01, eps=0.0,
                                            edge_feats=ef)
                a_e = attention_weights(plan, saved_e, H, 0.0)
            e_out = self.fc_edge_src(efeats_e).view(-1, H, De) * a_e.unsqueeze(-1)   # :295-299
            rets.append(e_out)
        if get_attention:
            rets.append(attention_weights(plan, saved, H, 0.0).unsqueeze(-1))
        return tuple(rets)


class EGATConv(EEGATConv):
    """``dgl.nn.EGATConv`` signature: no ``edge_scoring`` kwarg, no ``efeats_e`` positional."""

    def __init__(self, in_node_feats, in_edge_feats, out_node_feats, out_edge_feats, num_heads,
                 bias=True):
        super().__init__(in_node_feats, in_edge_feats, out_node_feats, out_edge_feats, num_heads,
                         bias=bias, edge_scoring=False)

    def forward(self, graph, nfeats, efeats, edge_weight=None, get_attention=False):
        return super().forward(graph, nfeats, efeats, None, edge_weight, get_attention)
