"""Dense projection dispatch (``y = x @ W.T + b``) for the node/edge embedding MLPs and the
attention-layer projections."""
import os

import torch


def gemm_backend() -> str:
    return os.environ.get("WHEATLEY_B200_GEMM", "cublas")


def dense_linear(x, weight, bias=None):
    return torch.nn.functional.linear(x, weight, bias)
