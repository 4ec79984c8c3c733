"""Minimal PyG stand-in for generic/gatv3_conv.py (MessagePassing.propagate with
sum aggregation, dense Linear with glorot init, softmax, self-loop utils)."""
from . import nn, typing, utils  # noqa: F401
