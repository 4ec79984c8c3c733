import torch
from . import inits


class Linear(torch.nn.Linear):
    """torch_geometric.nn.dense.Linear with weight_initializer='glorot' and zero bias."""

    def __init__(self, in_channels, out_channels, bias=True, weight_initializer=None,
                 bias_initializer=None):
        self.weight_initializer = weight_initializer
        super().__init__(in_channels, out_channels, bias=bias)

    def reset_parameters(self):
        if self.weight_initializer == "glorot":
            inits.glorot(self.weight)
        else:
            super().reset_parameters()
            return
        if self.bias is not None:
            inits.zeros(self.bias)
