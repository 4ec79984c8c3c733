"""Factored edge features.

Both reference trunks feed the attention layer edge features that are a function of a
small integer edge *type* (jssp/models/gnn_dgl.py:364 ``edge_embedder[0](g.edata["type"])``;
psp/models/edge_embedder.py:130-143 ``Embedding(type) + Linear(att_rc) + Linear(att_rp)``).
Passing the factors instead of the materialised ``[E, F]`` tensor lets the layer project
the T-row table once and gather inside the kernel (no ``[E,F]x[F,H*D]`` GEMM, no
``[E,H*D]`` intermediate)."""
import torch
from torch import nn


class TypedEdgeFeats:
    """``efeats[e] = weight[etype[e]]`` - drop-in for the dense ``[E, F]`` tensor."""

    def __init__(self, etype: torch.Tensor, weight: torch.Tensor):
        self.etype, self.weight = etype, weight
        self.attrs = None
        self.attr_weight = None
        self.tables = None          # optional {id(layer weight): projected [T, H*D] table} (edge_tables)

    @property
    def shape(self):
        return (self.etype.shape[0], self.weight.shape[1])

    def dense(self) -> torch.Tensor:
        return torch.nn.functional.embedding(self.etype.long(), self.weight)


class _EdgeTables(torch.autograd.Function):
    """Projected edge-type tables of ALL layers at once: ``table_l = weight @ W_l.T + b_l`` ([T, H*D] each).
    Per layer this is ``fc_fij(Embedding.weight) + bias`` (jssp/models/gnn_dgl.py:364 + generic/eegatconv.py:262,
    271-275 on the T distinct rows); layer by layer it costs one 6-row GEMM forward and two backward plus a bias
    sum - 30 launches of ~3 us of GPU and ~8 us of host time each for 6 layers.  Here: 4 launches forward, 4
    backward, same arithmetic (one GEMM over the concatenated weights)."""

    @staticmethod
    def forward(ctx, weight, n_layers, *params):
        Ws, bs = params[:n_layers], params[n_layers:]
        T, HD = weight.shape[0], Ws[0].shape[0]
        Wcat = torch.cat(Ws, 0)                                           # [L*HD, F]
        if bs[0] is not None:
            out = torch.addmm(torch.cat(bs, 0), weight, Wcat.t())         # [T, L*HD]
        else:
            out = weight @ Wcat.t()
        tabs = out.view(T, n_layers, HD).transpose(0, 1).contiguous()     # [L, T, HD]
        ctx.save_for_backward(weight, Wcat)
        ctx.n_layers, ctx.has_bias, ctx.HD = n_layers, bs[0] is not None, HD
        ctx.set_materialize_grads(False)
        return tabs.unbind(0)

    @staticmethod
    def backward(ctx, *gs):
        weight, Wcat = ctx.saved_tensors
        L, HD = ctx.n_layers, ctx.HD
        if all(g is None for g in gs):
            return (None,) * (2 + 2 * L)
        gs = [g if g is not None else torch.zeros((weight.shape[0], HD), dtype=weight.dtype, device=weight.device)
              for g in gs]
        g2 = torch.cat(gs, 1)                                             # [T, L*HD]
        need = ctx.needs_input_grad
        d_weight = g2 @ Wcat if need[0] else None                         # [T, F]
        dWcat = g2.t() @ weight                                           # [L*HD, F]
        dWs = tuple(dWcat[l * HD:(l + 1) * HD] if need[2 + l] else None for l in range(L))
        if ctx.has_bias:
            dbcat = g2.sum(0)
            dbs = tuple(dbcat[l * HD:(l + 1) * HD] if need[2 + L + l] else None for l in range(L))
        else:
            dbs = (None,) * L
        return (d_weight, None) + dWs + dbs


def edge_tables(weight, layer_weights, layer_biases):
    """``{id(W_l): table_l}`` for ``TypedEdgeFeats.tables`` (looked up by ``functional.project_edge_feats``)."""
    L = len(layer_weights)
    has_b = [b is not None for b in layer_biases]
    if L == 0 or (any(has_b) and not all(has_b)) or len({tuple(w.shape) for w in layer_weights}) != 1:
        return None
    tabs = _EdgeTables.apply(weight, L, *layer_weights, *layer_biases)
    return {id(w): t for w, t in zip(layer_weights, tabs)}


class AffineEdgeFeats(TypedEdgeFeats):
    """``efeats[e] = weight[etype[e]] + attrs[e] @ attr_weight.T``  with ``attrs [E, K]``,
    ``attr_weight [F, K]`` (K small: 5 for the PSP embedder, psp/models/edge_embedder.py:130-143,
    whose two ``Linear`` biases are folded into ``weight``)."""

    def __init__(self, etype, weight, attrs, attr_weight):
        super().__init__(etype, weight)
        self.attrs, self.attr_weight = attrs, attr_weight

    def dense(self) -> torch.Tensor:
        return super().dense() + self.attrs.to(self.attr_weight.dtype) @ self.attr_weight.t()


class PspEdgeEmbedder(nn.Module):
    """``edge_embedding_flavor == "sum"`` branch of psp/models/edge_embedder.py:25-143 with the
    reference parameter names (``edge_type_embedder``, ``rc_att_embedder``, ``rp_att_embedder``).

    ``forward(g)`` returns the factored :class:`AffineEdgeFeats`; ``.dense()`` of it is exactly
    the reference's ``[E, hidden]`` tensor (golden: tests/golden/edge_embedder.npz)."""

    def __init__(self, edge_embedding_flavor, max_n_resources, hidden_dim, add_rp_edges,
                 factored_rp, n_mlp_layers_features_extractor=None,
                 activation_features_extractor=None):
        super().__init__()
        if edge_embedding_flavor != "sum":
            raise NotImplementedError("only the default 'sum' edge embedding is on the live path "
                                      "(args.py default; SURVEY F13)")
        self.edge_embedding_flavor = edge_embedding_flavor
        self.max_n_resources = max_n_resources
        self.hidden_dim = hidden_dim
        self.add_rp_edges = add_rp_edges
        self.factored_rp = factored_rp
        self.n_edge_type = 16
        self.edge_type_embedder = nn.Embedding(self.n_edge_type, hidden_dim)
        self.rc_att_embedder = nn.Linear(2, hidden_dim)
        if add_rp_edges != "none":
            self.rp_att_embedder = nn.Linear(3 * max_n_resources if factored_rp else 3, hidden_dim)

    def factor(self, etype, att_rc=None, att_rp=None) -> AffineEdgeFeats:
        table = self.edge_type_embedder.weight
        attrs, ws = [], []
        if att_rc is not None:                      # psp/models/edge_embedder.py:134-137
            table = table + self.rc_att_embedder.bias
            attrs.append(att_rc.float())
            ws.append(self.rc_att_embedder.weight)
        if self.add_rp_edges != "none" and att_rp is not None:   # :138-142
            table = table + self.rp_att_embedder.bias
            attrs.append(att_rp.float())
            ws.append(self.rp_att_embedder.weight)
        if not attrs:
            return TypedEdgeFeats(etype, table)
        return AffineEdgeFeats(etype, table, torch.cat(attrs, dim=1), torch.cat(ws, dim=1))

    def forward(self, g):
        def get(key):
            try:
                return g.edata(None, key)
            except KeyError:
                return None
        a5 = get("attr5")       # the device batch builder's [E, 5] = att_rc | att_rp (one tensor for every layer)
        if a5 is not None and self.add_rp_edges != "none":
            table = self.edge_type_embedder.weight + self.rc_att_embedder.bias + self.rp_att_embedder.bias
            return AffineEdgeFeats(g.edata(None, "type"), table, a5,
                                   torch.cat([self.rc_att_embedder.weight, self.rp_att_embedder.weight], dim=1))
        return self.factor(g.edata(None, "type"), get("att_rc"), get("att_rp"))
