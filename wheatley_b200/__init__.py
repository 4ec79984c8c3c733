"""wheatley_b200 - B200-native (sm_100a) graph-attention hot path for Wheatley.

Host-side mirror of the reference's module interfaces for this one path
(SURVEY.md section 8b); all arithmetic runs in ``libb200gat.so`` (include/b200gat.h).
Importing the package loads the library and fails loudly if it is missing - there is no
CPU / PyTorch fallback.

================  ==========================================================
reference         drop-in here
================  ==========================================================
dgl.nn.EGATConv   wheatley_b200.EGATConv          (jssp/models/gnn_dgl.py:30)
EEGATConv         wheatley_b200.EEGATConv         (generic/eegatconv.py:18)
GraphConv         wheatley_b200.GraphConv         (psp/graph/graph_conv.py:8)
GATv2Conv (PyG)   wheatley_b200.GATv2Conv         (psp/graph/graph_conv.py:33)
GATv3Conv         wheatley_b200.GATv3Conv         (generic/gatv3_conv.py:16)
MLP               wheatley_b200.MLP               (generic/mlp.py:29)
GnnMP / GnnFlat   wheatley_b200.GnnMP / GnnFlat   (psp/models/gnn_mp.py:38, gnn_flat.py:7)
Agent.preprocess  wheatley_b200.psp_batch         (psp/models/agent.py:298-406)
obs wire format   wheatley_b200.wire              (psp/graph/pyg_graph.py:240-283)
Agent head        wheatley_b200.agent_head        (generic/agent.py:190-277)
================  ==========================================================
"""
from . import _lib  # noqa: F401  (loads libb200gat.so or raises ImportError)
from ._lib import B200GatError
from .plan import GraphPlan, plan_for_edge_index
from .edge_feats import TypedEdgeFeats, AffineEdgeFeats
from .functional import (EdgeFeatures, fused_gat_conv, attention_weights, edge_output,
                         masked_log_softmax, masked_logprob_entropy)
from .eegatconv import EGATConv, EEGATConv, DGLError
from .graph_conv import GraphConv, GATv2Conv, GATv3Conv
from .mlp import MLP
from .gnn_mp import GnnFlat, GnnMP
from . import agent_head
from .gnn_dgl import GnnDGL
from .agent import JsspAgent, PspAgent

__all__ = [
    "B200GatError", "GraphPlan", "plan_for_edge_index", "TypedEdgeFeats", "AffineEdgeFeats",
    "EdgeFeatures", "fused_gat_conv", "attention_weights", "edge_output", "masked_log_softmax",
    "masked_logprob_entropy", "EGATConv", "EEGATConv", "DGLError", "GraphConv", "GATv2Conv",
    "GATv3Conv", "MLP", "GnnMP", "GnnFlat", "GnnDGL", "JsspAgent", "PspAgent", "agent_head",
]
