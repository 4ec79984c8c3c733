"""GraphConv / GATv2Conv / GATv3Conv: B200-native drop-ins for the PyG-flavour boundary
(psp/graph/graph_conv.py:8-60, ``torch_geometric.nn.conv.GATv2Conv`` as configured at
psp/graph/graph_conv.py:33-40, generic/gatv3_conv.py:16-213).

Same constructors, ``forward`` signatures, parameter names (``conv.lin_l.weight`` ...,
so reference checkpoints load, SURVEY section 5 "Checkpoint") and initialisation; the
``propagate`` body (gather, leaky_relu, attention dot, ``softmax`` by target, scatter-sum)
is one fused CUDA kernel with a fused backward (include/b200gat.h)."""
import math

import torch
from torch import nn

from . import _lib
from .dense import dense_linear
from .edge_feats import TypedEdgeFeats
from .eegatconv import EEGATConv
from .functional import attention_weights, fused_gat_conv, project_edge_feats
from .plan import GraphPlan, plan_for_edge_index


def _glorot(t):
    """torch_geometric.nn.inits.glorot."""
    a = math.sqrt(6.0 / (t.size(-2) + t.size(-1)))
    with torch.no_grad():
        t.uniform_(-a, a)


class _PygLinear(nn.Linear):
    """``torch_geometric.nn.dense.Linear(weight_initializer='glorot')``: glorot weight, zero bias."""

    def reset_parameters(self):
        _glorot(self.weight)
        if self.bias is not None:
            nn.init.zeros_(self.bias)

    def forward(self, x):
        return dense_linear(x, self.weight, self.bias)


def _resolve(edge_index, n_nodes):
    if isinstance(edge_index, GraphPlan):
        return edge_index
    plan = getattr(edge_index, "plan", None)
    if isinstance(plan, GraphPlan):
        return plan
    return plan_for_edge_index(edge_index, n_nodes)


class GATv2Conv(nn.Module):
    """PyG ``GATv2Conv(in, out, heads, add_self_loops=False, edge_dim=..., bias=True)`` with
    ``concat=True, negative_slope=0.2, dropout=0, share_weights=False`` (the only configuration
    Wheatley builds, psp/graph/graph_conv.py:33-40).

    ``forward(x, edge_index, edge_attr)``: ``edge_index [2,E]`` int64 (row 0 = source, row 1 =
    target) or a prebuilt :class:`GraphPlan`; ``edge_attr`` a dense ``[E, edge_dim]`` tensor, a
    :class:`TypedEdgeFeats` / ``AffineEdgeFeats`` factorisation, or ``None``
    (psp/models/gnn_hier.py:69,115)."""

    def __init__(self, in_channels, out_channels, heads=1, concat=True, negative_slope=0.2,
                 dropout=0.0, add_self_loops=False, edge_dim=None, fill_value="mean", bias=True,
                 share_weights=False, **kwargs):
        super().__init__()
        if not concat or dropout != 0.0 or add_self_loops or share_weights:
            raise NotImplementedError("Wheatley builds GATv2Conv with concat=True, dropout=0, "
                                      "add_self_loops=False, share_weights=False "
                                      "(psp/graph/graph_conv.py:33-40)")
        self.in_channels, self.out_channels, self.heads = in_channels, out_channels, heads
        self.negative_slope, self.edge_dim = negative_slope, edge_dim
        self.lin_l = _PygLinear(in_channels, heads * out_channels, bias=bias)
        self.lin_r = _PygLinear(in_channels, heads * out_channels, bias=bias)
        self.att = nn.Parameter(torch.empty(1, heads, out_channels))
        self.lin_edge = _PygLinear(edge_dim, heads * out_channels, bias=False) if edge_dim is not None else None
        if bias:
            self.bias = nn.Parameter(torch.empty(heads * out_channels))
        else:
            self.register_parameter("bias", None)
        self.reset_parameters()

    def reset_parameters(self):
        self.lin_l.reset_parameters()
        self.lin_r.reset_parameters()
        if self.lin_edge is not None:
            self.lin_edge.reset_parameters()
        _glorot(self.att)
        if self.bias is not None:
            nn.init.zeros_(self.bias)

    def forward(self, x, edge_index, edge_attr=None, return_attention_weights=None):
        plan = _resolve(edge_index, x.shape[0])
        x_l = self.lin_l(x)                     # source term of the logit AND the aggregated value
        x_r = self.lin_r(x)
        ef = None
        if edge_attr is not None:
            if self.lin_edge is None:
                raise _lib.B200GatError("edge_attr given but the layer was built with edge_dim=None")
            ef = project_edge_feats(plan, edge_attr, self.lin_edge.weight, None)
        out, saved = fused_gat_conv(plan, x_l, x_r, None, self.att, heads=self.heads,
                                    dim=self.out_channels, flavour=_lib.FLAVOUR_ADDITIVE,
                                    slope=self.negative_slope, eps=1e-16, out_bias=self.bias,
                                    edge_feats=ef)
        if isinstance(return_attention_weights, bool):
            return out, (edge_index, attention_weights(plan, saved, self.heads, 1e-16))
        return out

    def __repr__(self):
        return f"{type(self).__name__}({self.in_channels}, {self.out_channels}, heads={self.heads})"


class GATv3Conv(nn.Module):
    """generic/gatv3_conv.py:16-213.  ``s = Linear(C->1)(x_i * x_j) / sqrt(C)`` shared by all
    heads (:60-64,186-192); self loops are removed then re-added (:149-153) - done on the
    device with torch index ops, then planned like any other graph.  ``edge_attr`` /
    ``lin_edge`` are accepted and unused, as in the reference's ``message`` (:182).  Reference
    quirk kept: with ``concat=True`` the layer ends up with NO output bias (:108-113)."""

    def __init__(self, in_channels, out_channels, heads=1, dropout=0.0, edge_dim=None,
                 fill_value="mean", name="unnamed", *, concat=True, bias=True, share_weights=False,
                 **kwargs):
        super().__init__()
        if not concat or dropout != 0.0 or share_weights or not isinstance(in_channels, int):
            raise NotImplementedError("Wheatley builds GATv3Conv with concat=True, dropout=0 "
                                      "(psp/graph/graph_conv.py:24-31)")
        self.in_channels, self.out_channels, self.heads = in_channels, out_channels, heads
        self.edge_dim, self.name = edge_dim, name
        self.context_attention_layer = nn.Linear(out_channels, 1, bias=True)
        self.lin_l = _PygLinear(in_channels, heads * out_channels, bias=bias)
        self.lin_r = _PygLinear(in_channels, heads * out_channels, bias=bias)
        self.lin_edge = _PygLinear(edge_dim, heads * out_channels, bias=False) if edge_dim is not None else None
        self.register_parameter("bias", None)           # generic/gatv3_conv.py:108-113
        self._loop_cache = None

    def _with_self_loops(self, edge_index, n):
        tag = (edge_index._version, tuple(edge_index.shape), n)
        if (self._loop_cache is not None and self._loop_cache[2] is edge_index
                and self._loop_cache[0] == tag):            # the same tensor object, unchanged
            return self._loop_cache[1]
        keep = edge_index[0] != edge_index[1]            # remove_self_loops keeps edge order
        loop = torch.arange(n, dtype=edge_index.dtype, device=edge_index.device)
        ei = torch.cat([edge_index[:, keep], torch.stack([loop, loop])], dim=1)   # add_self_loops appends
        self._loop_cache = (tag, ei, edge_index)
        return ei

    def forward(self, x, edge_index, edge_attr=None, *, return_attention_weights=None):
        if isinstance(edge_index, GraphPlan):
            raise _lib.B200GatError("GATv3Conv rewrites the self loops: pass edge_index, not a plan")
        ei = self._with_self_loops(edge_index, x.shape[0])
        plan = plan_for_edge_index(ei, x.shape[0])
        x_l = self.lin_l(x)
        x_r = self.lin_r(x)
        out, saved = fused_gat_conv(plan, x_l, x_r, None, self.context_attention_layer.weight,
                                    heads=self.heads, dim=self.out_channels, flavour=_lib.FLAVOUR_DOT,
                                    eps=1e-16, attn_bias=self.context_attention_layer.bias,
                                    out_bias=self.bias)
        if isinstance(return_attention_weights, bool):
            return out, (ei, attention_weights(plan, saved, self.heads, 1e-16))
        return out

    def __repr__(self):
        return f"GATv3Conv({self.in_channels}, {self.out_channels}, heads={self.heads})"


class GraphConv(nn.Module):
    """psp/graph/graph_conv.py:8-60 - same constructor, attributes (``gatv3``, ``pyg``,
    ``conv``) and return conventions."""

    def __init__(self, in_dim, out_dim, num_heads, bias=True, edge_scoring=False, pyg=False,
                 edge_dim=None):
        super().__init__()
        self.gatv3 = False                                # psp/graph/graph_conv.py:20
        self.pyg = pyg
        if self.pyg:
            if self.gatv3:
                self.conv = GATv3Conv(in_dim, out_dim, num_heads, dropout=0,
                                      edge_dim=in_dim if edge_dim is None else edge_dim, bias=bias)
            else:
                self.conv = GATv2Conv(in_dim, out_dim, num_heads, add_self_loops=False,
                                      edge_dim=in_dim if edge_dim is None else edge_dim, bias=bias)
        else:
            self.conv = EEGATConv(in_dim, out_dim, in_dim, out_dim, num_heads=num_heads,
                                  edge_scoring=edge_scoring)

    def forward(self, g, node_feats, edge_feats, edge_efeats=None):
        if self.pyg:
            plan = getattr(g, "plan", None)
            use_plan = isinstance(plan, GraphPlan) and not self.gatv3
            return self.conv(node_feats, plan if use_plan else g.edge_index, edge_feats), None
        f, ef = self.conv(g, node_feats, edge_feats, edge_efeats)[:2]
        return f.flatten(start_dim=-2, end_dim=-1), ef

    def forward_nog(self, node_feats, edge_index, edge_feats, edge_efeats=None):
        return self.conv(node_feats, edge_index, edge_feats), None
