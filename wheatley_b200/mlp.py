"""MLP - drop-in for generic/mlp.py:29-77 (same constructor, ``layers`` / ``norms`` /
``activation_layer`` attribute and parameter names, so ``state_dict`` keys match).

Dense projections are the one GEMM-shaped part of the path (north_star item 4).
``dense_linear`` dispatches each ``Linear`` to the library's tcgen05 TF32 GEMM
(``b200gat_linear_tf32_*``, fp32 in / fp32 accumulate in TMEM / fp32 out - the precision the
reference runs its matmuls in, ``torch.set_float32_matmul_precision("high")``,
psp/train_psp.py:47) when the shape is covered and ``WHEATLEY_B200_GEMM`` is not ``cublas``;
otherwise to cuBLAS through ``torch.nn.functional.linear`` (a plain library GEMM)."""
import torch
from torch import nn

from .dense import dense_linear

_ACT = {
    "tanh": nn.Tanh,
    "relu": nn.LeakyReLU,          # sic: generic/mlp.py:61 maps "relu" to LeakyReLU
    "elu": nn.ELU,
    "gelu": nn.GELU,
    "selu": nn.SELU,
    "silu": nn.SiLU,
}


class MLP(nn.Module):
    def __init__(self, n_layers, input_dim, hidden_dim, output_dim, norm, activation):
        super().__init__()
        if n_layers < 1:
            raise Exception("Number of layers must be >= 1")
        self.n_layers = n_layers
        self.layers = nn.ModuleList()
        self.norm = norm
        if self.norm:
            self.norms = nn.ModuleList()
        last_dim = input_dim
        for _ in range(self.n_layers - 1):
            self.layers.append(nn.Linear(last_dim, hidden_dim))
            if self.norm:
                self.norms.append(nn.LayerNorm(hidden_dim))
            last_dim = hidden_dim
        self.layers.append(nn.Linear(last_dim, output_dim))
        if isinstance(activation, str):
            activation = _ACT[activation]
        self.activation_layer = activation()

    def forward(self, x):
        for layer in range(self.n_layers - 1):
            x = dense_linear(x, self.layers[layer].weight, self.layers[layer].bias)
            if self.norm:
                x = self.norms[layer](x)
            x = self.activation_layer(x)
        return dense_linear(x, self.layers[-1].weight, self.layers[-1].bias)
