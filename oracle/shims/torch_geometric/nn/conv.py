import torch


class MessagePassing(torch.nn.Module):
    """propagate() for flow='source_to_target', node_dim=0, aggr='sum' with x given
    as a pair: x_j = x[0][edge_index[0]], x_i = x[1][edge_index[1]]."""

    def __init__(self, aggr="sum", flow="source_to_target", node_dim=0, **kwargs):
        super().__init__()
        assert aggr == "sum" and flow == "source_to_target" and node_dim == 0
        self.aggr, self.flow, self.node_dim = aggr, flow, node_dim

    def propagate(self, edge_index, size=None, **kwargs):
        x = kwargs["x"]
        x_j = x[0].index_select(0, edge_index[0])
        x_i = x[1].index_select(0, edge_index[1])
        msg = self.message(x_j=x_j, x_i=x_i, edge_index=edge_index)
        out = self.aggregate(msg, edge_index[1], ptr=None, dim_size=size[1])
        return self.update(out)

    def aggregate(self, inputs, index, ptr=None, dim_size=None):
        out = torch.zeros((dim_size,) + tuple(inputs.shape[1:]), dtype=inputs.dtype)
        return out.index_add(0, index, inputs)

    def update(self, inputs):
        return inputs


class GATv2Conv(MessagePassing):
    """Restatement of ``torch_geometric.nn.conv.GATv2Conv`` (PyG 2.5 / 2.6 series,
    ``torch_geometric/nn/conv/gatv2_conv.py``: ``__init__``, ``reset_parameters``, ``forward``,
    ``edge_update``, ``message``) for the arguments Wheatley passes
    (psp/graph/graph_conv.py:33-40: ``GATv2Conv(in_dim, out_dim, num_heads, add_self_loops=False,
    edge_dim=..., bias=bias)``; everything else at its default: ``concat=True, negative_slope=0.2,
    dropout=0.0, fill_value='mean', share_weights=False``).

    PARITY UNPINNED (third-party layer): PyG is un-vendored, its version is not pinned by the reference
    (requirements.txt:16) and it cannot be installed here, so this class is written from the published
    source, not executed from it.  What it fixes, line by line of that source:
      * ``lin_l`` / ``lin_r``: ``Linear(in, heads*out, bias=bias, weight_initializer='glorot')`` (zero bias init),
        ``lin_edge``: ``Linear(edge_dim, heads*out, bias=False, weight_initializer='glorot')``,
        ``att``: ``Parameter(1, heads, out)`` glorot, ``bias``: ``Parameter(heads*out)`` zeros (concat);
      * ``forward``: ``x_l = lin_l(x).view(-1,H,C)``, ``x_r = lin_r(x).view(-1,H,C)``; no self-loop rewrite
        (``add_self_loops=False``); ``alpha = edge_updater(edge_index, x=(x_l, x_r), edge_attr)``;
        ``out = propagate(edge_index, x=(x_l, x_r), alpha=alpha)``; ``out.view(-1, H*C) + bias``;
      * ``edge_update``: ``x = x_i + x_j`` (i = target = edge_index[1], j = source = edge_index[0]);
        ``+ lin_edge(edge_attr).view(-1,H,C)``; ``leaky_relu(x, 0.2)``; ``alpha = (x * att).sum(-1)``;
        ``softmax(alpha, index=edge_index[1], num_nodes=N)`` (max-subtracted, ``+1e-16``); dropout p=0;
      * ``message``: ``x_j * alpha.unsqueeze(-1)``, summed at the target node."""

    def __init__(self, in_channels, out_channels, heads=1, concat=True, negative_slope=0.2, dropout=0.0,
                 add_self_loops=True, edge_dim=None, fill_value="mean", bias=True, share_weights=False,
                 **kwargs):
        super().__init__(node_dim=0, **kwargs)
        from .dense import Linear
        from . import inits
        assert concat and dropout == 0.0 and not add_self_loops and not share_weights
        self.in_channels, self.out_channels, self.heads = in_channels, out_channels, heads
        self.concat, self.negative_slope, self.dropout = concat, negative_slope, dropout
        self.add_self_loops, self.edge_dim, self.fill_value = add_self_loops, edge_dim, fill_value
        self.share_weights = share_weights
        self.lin_l = Linear(in_channels, heads * out_channels, bias=bias, weight_initializer="glorot")
        self.lin_r = Linear(in_channels, heads * out_channels, bias=bias, weight_initializer="glorot")
        self.att = torch.nn.Parameter(torch.empty(1, heads, out_channels))
        if edge_dim is not None:
            self.lin_edge = Linear(edge_dim, heads * out_channels, bias=False, weight_initializer="glorot")
        else:
            self.lin_edge = None
        if bias:
            self.bias = torch.nn.Parameter(torch.empty(heads * out_channels))
        else:
            self.register_parameter("bias", None)
        self.lin_l.reset_parameters()
        self.lin_r.reset_parameters()
        if self.lin_edge is not None:
            self.lin_edge.reset_parameters()
        inits.glorot(self.att)
        inits.zeros(self.bias)

    def forward(self, x, edge_index, edge_attr=None, return_attention_weights=None):
        from ..utils import softmax
        H, C = self.heads, self.out_channels
        x_l = self.lin_l(x).view(-1, H, C)
        x_r = self.lin_r(x).view(-1, H, C)
        x_j = x_l.index_select(0, edge_index[0])
        x_i = x_r.index_select(0, edge_index[1])
        z = x_i + x_j                                              # edge_update
        if edge_attr is not None:
            if edge_attr.dim() == 1:
                edge_attr = edge_attr.view(-1, 1)
            z = z + self.lin_edge(edge_attr).view(-1, H, C)
        z = torch.nn.functional.leaky_relu(z, self.negative_slope)
        alpha = (z * self.att).sum(dim=-1)
        alpha = softmax(alpha, edge_index[1], None, x.shape[0])
        msg = x_j * alpha.unsqueeze(-1)                            # message
        out = self.aggregate(msg, edge_index[1], ptr=None, dim_size=x.shape[0])
        out = out.view(-1, H * C)
        if self.bias is not None:
            out = out + self.bias
        if isinstance(return_attention_weights, bool):
            return out, (edge_index, alpha)
        return out
