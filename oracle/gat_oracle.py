"""CPU restatement (plain PyTorch, any float dtype) of the three graph-attention
layer flavours on Wheatley's hot path.  TEST INFRASTRUCTURE - see oracle/__init__.py.

All ``file:line`` citations are relative to the reference tree
(jolibrain/wheatley @ 7143392).

Conventions: ``src``/``dst`` are int64 ``[E]`` tensors, edge ``e`` goes
``src[e] -> dst[e]`` and messages are reduced at ``dst`` (DGL ``update_all`` /
PyG ``flow="source_to_target"``).
"""
from __future__ import annotations

import math

import torch
from torch import nn


# --------------------------------------------------------------------------
# primitives
# --------------------------------------------------------------------------
def segment_softmax(score: torch.Tensor, dst: torch.Tensor, n_nodes: int, eps: float = 0.0):
    """Softmax of ``score [E, H]`` over the incoming edges of every destination.

    DGL ``edge_softmax`` (norm_by="dst", generic/eegatconv.py:280):
    ``exp(s - max_dst) / sum_dst`` with ``eps = 0``.
    PyG ``torch_geometric.utils.softmax`` (generic/gatv3_conv.py:193):
    same with the max detached and ``eps = 1e-16`` added to the denominator.
    """
    E, H = score.shape
    idx = dst.view(-1, 1).expand(E, H)
    smax = torch.full((n_nodes, H), -math.inf, dtype=score.dtype)
    smax = smax.scatter_reduce(0, idx, score.detach(), reduce="amax", include_self=True)
    ex = torch.exp(score - smax.index_select(0, dst))
    ssum = torch.zeros((n_nodes, H), dtype=score.dtype).index_add(0, dst, ex)
    return ex / (ssum.index_select(0, dst) + eps)


def aggregate(alpha: torch.Tensor, value: torch.Tensor, src, dst, n_nodes: int):
    """``out[v] = sum_{e->v} alpha[e,h] * value[src_e,h,:]``
    (DGL ``update_all(u_mul_e, sum)`` generic/eegatconv.py:292; PyG scatter-sum
    generic/gatv3_conv.py:198-207)."""
    msg = value.index_select(0, src) * alpha.unsqueeze(-1)
    out = torch.zeros((n_nodes,) + tuple(value.shape[1:]), dtype=value.dtype)
    return out.index_add(0, dst, msg)


# --------------------------------------------------------------------------
# flavour A : EGAT / EEGAT   (generic/eegatconv.py:188-313 == dgl.nn.EGATConv)
# --------------------------------------------------------------------------
def egat_forward(x, src, dst, efeats, W_ni, W_nj, W_fij, W_src, attn, bias,
                 edge_weight=None, slope: float = 0.01):
    """Returns ``(h_out [N,H,Do], f_out [E,H,De], a [E,H,1])``.

    x [N,Fin]; efeats [E,Fe]; W_ni/W_nj [H*De,Fin]; W_fij [H*De,Fe];
    W_src [H*Do,Fin]; attn [1,H,De]; bias [H*De] or None.
    """
    N = x.shape[0]
    H, De = attn.shape[1], attn.shape[2]
    if bool((torch.bincount(dst, minlength=N) == 0).any()):     # eegatconv.py:238-247
        raise ValueError("There are 0-in-degree nodes in the graph")
    f_ni = x @ W_ni.t()                                           # :260
    f_nj = x @ W_nj.t()                                           # :261
    f_fij = efeats @ W_fij.t()                                    # :262
    f_out = f_ni.index_select(0, src) + f_nj.index_select(0, dst) # :266-269  u_add_v
    f_out = f_out + f_fij                                         # :271
    if bias is not None:
        f_out = f_out + bias                                      # :274-275
    f_out = nn.functional.leaky_relu(f_out, slope)                # :276 (torch default 0.01)
    f_out = f_out.view(-1, H, De)                                 # :277
    e = (f_out * attn).sum(dim=-1)                                # :279  [E,H]
    a = segment_softmax(e, dst, N, eps=0.0)                       # :280
    if edge_weight is not None:                                   # :284-287
        a = a * edge_weight.view(-1, 1)
    h = (x @ W_src.t()).view(N, H, -1)                            # :288-290
    h_out = aggregate(a, h, src, dst, N)                          # :292
    return h_out, f_out, a.unsqueeze(-1)


def eegat_forward_edge_scoring(x, src, dst, efeats, efeats_e, p, slope: float = 0.01):
    """``edge_scoring=True`` branch of EEGATConv (generic/eegatconv.py:263-264,
    272-273,281-283,294-299).  ``p`` is a dict of weights keyed by the reference
    parameter names.  Returns ``(h_out, f_out, e_out, a)``."""
    N = x.shape[0]
    attn = p["attn"]
    H, De = attn.shape[1], attn.shape[2]
    if efeats_e is None:
        efeats_e = efeats                                         # :257-258
    f_ni = x @ p["fc_ni.weight"].t()
    f_nj = x @ p["fc_nj.weight"].t()
    f_fij = efeats @ p["fc_fij.weight"].t()
    f_fij_e = efeats_e @ p["fc_fij_e.weight"].t()                 # :263-264
    f_out = f_ni.index_select(0, src) + f_nj.index_select(0, dst) + f_fij
    f_out = f_out + f_fij_e                                       # :272-273
    if p.get("bias") is not None:
        f_out = f_out + p["bias"]
    f_out = nn.functional.leaky_relu(f_out, slope).view(-1, H, De)
    a = segment_softmax((f_out * attn).sum(-1), dst, N)
    a_e = segment_softmax((f_out * p["attn_e"]).sum(-1), dst, N)  # :281-283
    h = (x @ p["fc_node_src.weight"].t()).view(N, H, -1)
    h_out = aggregate(a, h, src, dst, N)
    f_e = (efeats_e @ p["fc_edge_src.weight"].t()).view(-1, H, De)  # :295-297
    e_out = f_e * a_e.unsqueeze(-1)                               # :182-186, :298-299
    return h_out, f_out, e_out, a.unsqueeze(-1)


# --------------------------------------------------------------------------
# flavour B : PyG GATv2Conv as configured at psp/graph/graph_conv.py:33-40
#             (add_self_loops=False, edge_dim=in_dim, bias=True, concat=True,
#              share_weights=False, negative_slope=0.2, dropout=0)
# --------------------------------------------------------------------------
def gatv2_forward(x, edge_index, edge_attr, W_l, b_l, W_r, b_r, W_e, att, bias,
                  slope: float = 0.2):
    """Returns ``(out [N,H*C], alpha [E,H])``.  ``edge_attr`` may be None
    (psp/models/gnn_hier.py:69,115)."""
    src, dst = edge_index[0], edge_index[1]
    N = x.shape[0]
    H, C = att.shape[1], att.shape[2]
    x_l = x @ W_l.t()
    x_r = x @ W_r.t()
    if b_l is not None:
        x_l = x_l + b_l
    if b_r is not None:
        x_r = x_r + b_r
    x_l = x_l.view(N, H, C)
    x_r = x_r.view(N, H, C)
    z = x_l.index_select(0, src) + x_r.index_select(0, dst)      # x_j + x_i
    if edge_attr is not None:
        z = z + (edge_attr @ W_e.t()).view(-1, H, C)              # lin_edge has no bias
    z = nn.functional.leaky_relu(z, slope)
    s = (z * att).sum(dim=-1)                                     # [E,H]
    alpha = segment_softmax(s, dst, N, eps=1e-16)
    out = aggregate(alpha, x_l, src, dst, N).reshape(N, H * C)    # concat heads
    if bias is not None:
        out = out + bias
    return out, alpha


# --------------------------------------------------------------------------
# flavour C : GATv3Conv  (generic/gatv3_conv.py:122-198)
# --------------------------------------------------------------------------
def remove_then_add_self_loops(edge_index: torch.Tensor, n_nodes: int):
    """generic/gatv3_conv.py:149-153 - PyG ``remove_self_loops`` keeps edge
    order, ``add_self_loops`` appends ``(i,i)`` for i in range(N)."""
    keep = edge_index[0] != edge_index[1]
    ei = edge_index[:, keep]
    loop = torch.arange(n_nodes, dtype=edge_index.dtype)
    return torch.cat([ei, torch.stack([loop, loop])], dim=1)


def gatv3_forward(x, edge_index, W_l, b_l, W_r, b_r, w_c, b_c, bias, heads: int):
    """Returns ``(out [N,H*C], alpha [E',H], edge_index')`` where ``E'`` counts
    the edges after self-loop replacement.  ``w_c [1,C]``, ``b_c [1]`` are
    ``context_attention_layer`` (shared by all heads, :60-64)."""
    N = x.shape[0]
    x_l = x @ W_l.t()
    x_r = x @ W_r.t()
    if b_l is not None:
        x_l = x_l + b_l
    if b_r is not None:
        x_r = x_r + b_r
    C = x_l.shape[1] // heads
    x_l = x_l.view(N, heads, C)                                   # :135
    x_r = x_r.view(N, heads, C)                                   # :139
    ei = remove_then_add_self_loops(edge_index, N)                # :152-153
    src, dst = ei[0], ei[1]
    prod = x_r.index_select(0, dst) * x_l.index_select(0, src)    # :186  x_i * x_j
    s = (prod @ w_c.t()).squeeze(-1) + b_c                        # :187
    s = s / math.sqrt(C)                                          # :191-192
    alpha = segment_softmax(s, dst, N, eps=1e-16)                 # :193
    out = aggregate(alpha, x_l, src, dst, N).reshape(N, heads * C)  # :198, :167
    if bias is not None:
        out = out + bias                                          # :171-172
    return out, alpha, ei


# --------------------------------------------------------------------------
# nn.Module wrappers (reference parameter names; autograd gives the oracle
# gradients; run them in float64 for a tight reference)
# --------------------------------------------------------------------------
class OracleEGATConv(nn.Module):
    """Parameter names/shapes/init of generic/eegatconv.py:134-180 (edge_scoring=False)."""

    def __init__(self, in_node_feats, in_edge_feats, out_node_feats, out_edge_feats,
                 num_heads, bias=True):
        super().__init__()
        self.fc_node_src = nn.Linear(in_node_feats, out_node_feats * num_heads, bias=False)
        self.fc_ni = nn.Linear(in_node_feats, out_edge_feats * num_heads, bias=False)
        self.fc_nj = nn.Linear(in_node_feats, out_edge_feats * num_heads, bias=False)
        self.fc_fij = nn.Linear(in_edge_feats, out_edge_feats * num_heads, bias=False)
        self.attn = nn.Parameter(torch.empty(1, num_heads, out_edge_feats))
        if bias:
            self.bias = nn.Parameter(torch.empty(num_heads * out_edge_feats))
        else:
            self.register_buffer("bias", None)
        gain = nn.init.calculate_gain("relu")
        for w in (self.fc_node_src.weight, self.fc_ni.weight, self.fc_fij.weight,
                  self.fc_nj.weight, self.attn):
            nn.init.xavier_normal_(w, gain=gain)
        if bias:
            nn.init.constant_(self.bias, 0)

    def forward(self, src, dst, nfeats, efeats, edge_weight=None, get_attention=False):
        h, f, a = egat_forward(nfeats, src, dst, efeats, self.fc_ni.weight, self.fc_nj.weight,
                               self.fc_fij.weight, self.fc_node_src.weight, self.attn,
                               self.bias, edge_weight)
        return (h, f, a) if get_attention else (h, f)


def glorot_(t: torch.Tensor):
    """torch_geometric.nn.inits.glorot: U(-a, a), a = sqrt(6 / (size(-2)+size(-1)))."""
    a = math.sqrt(6.0 / (t.size(-2) + t.size(-1)))
    with torch.no_grad():
        t.uniform_(-a, a)
    return t


class OracleGATv2Conv(nn.Module):
    """PyG GATv2Conv parameter names: lin_l, lin_r, lin_edge, att, bias."""

    def __init__(self, in_channels, out_channels, heads, edge_dim=None, bias=True):
        super().__init__()
        self.heads, self.out_channels = heads, out_channels
        self.lin_l = nn.Linear(in_channels, heads * out_channels, bias=bias)
        self.lin_r = nn.Linear(in_channels, heads * out_channels, bias=bias)
        self.lin_edge = (nn.Linear(edge_dim, heads * out_channels, bias=False)
                         if edge_dim is not None else None)
        self.att = nn.Parameter(torch.empty(1, heads, out_channels))
        if bias:
            self.bias = nn.Parameter(torch.zeros(heads * out_channels))
        else:
            self.register_parameter("bias", None)
        glorot_(self.lin_l.weight)
        glorot_(self.lin_r.weight)
        if self.lin_edge is not None:
            glorot_(self.lin_edge.weight)
        glorot_(self.att)
        if bias:
            nn.init.zeros_(self.lin_l.bias)
            nn.init.zeros_(self.lin_r.bias)

    def forward(self, x, edge_index, edge_attr=None):
        out, _ = gatv2_forward(x, edge_index, edge_attr, self.lin_l.weight, self.lin_l.bias,
                               self.lin_r.weight, self.lin_r.bias,
                               None if self.lin_edge is None else self.lin_edge.weight,
                               self.att, self.bias)
        return out


class OracleGATv3Conv(nn.Module):
    """generic/gatv3_conv.py parameter names: context_attention_layer, lin_l, lin_r,
    lin_edge (unused in message).  Reference quirk (generic/gatv3_conv.py:108-113):
    with ``concat=True`` the ``else`` of the second ``if`` re-registers ``bias`` as
    None, so the layer never has an output bias and ``state_dict`` has no ``bias``
    key (confirmed by executing the reference file, tests/golden/make_golden.py)."""

    def __init__(self, in_channels, out_channels, heads, edge_dim=None, bias=True):
        super().__init__()
        self.heads, self.out_channels = heads, out_channels
        self.context_attention_layer = nn.Linear(out_channels, 1, bias=True)
        self.lin_l = nn.Linear(in_channels, heads * out_channels, bias=bias)
        self.lin_r = nn.Linear(in_channels, heads * out_channels, bias=bias)
        self.lin_edge = (nn.Linear(edge_dim, heads * out_channels, bias=False)
                         if edge_dim is not None else None)
        self.register_parameter("bias", None)
        glorot_(self.lin_l.weight)
        glorot_(self.lin_r.weight)
        if bias:
            nn.init.zeros_(self.lin_l.bias)
            nn.init.zeros_(self.lin_r.bias)
        if self.lin_edge is not None:
            glorot_(self.lin_edge.weight)

    def forward(self, x, edge_index, edge_attr=None):
        out, _, _ = gatv3_forward(x, edge_index, self.lin_l.weight, self.lin_l.bias,
                                  self.lin_r.weight, self.lin_r.bias,
                                  self.context_attention_layer.weight,
                                  self.context_attention_layer.bias, self.bias, self.heads)
        return out
