"""Factored edge features.

Both reference trunks feed the attention layer edge features that are a function of a
small integer edge *type* (jssp/models/gnn_dgl.py:364 ``edge_embedder[0](g.edata["type"])``;
psp/models/edge_embedder.py:130-143 ``Embedding(type) + Linear(att_rc) + Linear(att_rp)``).
Passing the factors instead of the materialised ``[E, F]`` tensor lets the layer project
the T-row table once and gather inside the kernel (no ``[E,F]x[F,H*D]`` GEMM, no
``[E,H*D]`` intermediate)."""
import torch


class TypedEdgeFeats:
    """``efeats[e] = weight[etype[e]]`` - drop-in for the dense ``[E, F]`` tensor."""

    def __init__(self, etype: torch.Tensor, weight: torch.Tensor):
        self.etype, self.weight = etype, weight

    @property
    def shape(self):
        return (self.etype.shape[0], self.weight.shape[1])

    def dense(self) -> torch.Tensor:
        return torch.nn.functional.embedding(self.etype.long(), self.weight)
