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
