"""MLP - drop-in for generic/mlp.py:29-77 (same constructor, ``layers`` / ``norms`` /
``activation_layer`` attribute and parameter names, so ``state_dict`` keys match).

Dense projections are the one GEMM-shaped part of the path (north_star item 4).
A GELU MLP without LayerNorm (the default configuration, args.py:335-348) runs as
``wheatley_b200.dense.gelu_mlp``: every ``Linear`` on the library's tcgen05 TF32 GEMM
(``b200gat_gemm_tf32``, fp32 in / fp32 accumulate in TMEM / fp32 out - the precision the reference
runs its matmuls in, ``torch.set_float32_matmul_precision("high")``, psp/train_psp.py:47) with bias,
GELU, GELU' and the residual add fused into the epilogues, when TF32 matmuls are allowed and
``WHEATLEY_B200_GEMM`` is not ``cublas``; otherwise cuBLAS through ``torch.nn.functional.linear``
(a plain library GEMM).  ``forward(x, residual=None)``: the optional ``residual`` (an extension; the
reference adds it right after the call, jssp/models/gnn_dgl.py:395-396) is fused into the last GEMM."""
import torch
from torch import nn

from .dense import dense_linear, gelu_mlp, gelu_mlp_applicable

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

    def forward(self, x, residual=None):
        if (not self.norm and isinstance(self.activation_layer, nn.GELU)
                and getattr(self.activation_layer, "approximate", "none") == "none"):
            x2 = x.reshape(-1, x.shape[-1])
            if gelu_mlp_applicable(x2, self.layers):
                r2 = None if residual is None else residual.reshape(-1, residual.shape[-1])
                y = gelu_mlp(x2, self.layers, r2)
                return y.reshape(*x.shape[:-1], y.shape[-1])
        y = self._forward_library(x)
        return y if residual is None else y + residual

    def _forward_library(self, x):
        for layer in range(self.n_layers - 1):
            x = dense_linear(x, self.layers[layer].weight, self.layers[layer].bias)
            if self.norm:
                x = self.norms[layer](x)
            x = self.activation_layer(x)
        return dense_linear(x, self.layers[-1].weight, self.layers[-1].bias)
