"""JsspAgent: ``gnn`` + ``value_net`` + ``action_net`` with the reference attribute names
(generic/agent.py:143-147, jssp/models/agent.py:70-121), so ``Agent.save``'s checkpoint layout
``{"gnn": ..., "value_net": ..., "action_net": ...}`` (generic/agent.py:158-172) round-trips.
The three head methods come from :class:`AgentHeadMixin` (generic/agent.py:190-277)."""
import torch
from torch import nn

from .agent_head import AgentHeadMixin
from .gnn_dgl import GnnDGL
from .mlp import MLP


class JsspAgent(AgentHeadMixin, nn.Module):
    def __init__(self, input_dim, max_n_nodes, max_n_machines, hidden_dim=64, n_layers=6, n_heads=4,
                 n_mlp_layers=3, activation="gelu", graph_pooling="learn", conflicts="clique",
                 hidden_dim_actor=64, hidden_dim_critic=64, n_mlp_layers_actor=3, n_mlp_layers_critic=3):
        nn.Module.__init__(self)
        self.gnn = GnnDGL(input_dim, "gatv2", graph_pooling, False, max_n_nodes, max_n_machines,
                          n_mlp_layers, n_layers, hidden_dim, activation, n_heads, residual=True,
                          normalize=False, conflicts=conflicts, split_output=True)
        feat_dim = input_dim + hidden_dim * (n_layers + 1)
        self.value_net = MLP(n_mlp_layers_critic, feat_dim, hidden_dim_critic, 1, False, activation)
        self.action_net = MLP(n_mlp_layers_actor, feat_dim, hidden_dim_actor, 1, False, activation)

    def save_dict(self):
        """generic/agent.py:158-172 layout (the two pickled specs are the caller's)."""
        return {"gnn": self.gnn.state_dict(), "value_net": self.value_net.state_dict(),
                "action_net": self.action_net.state_dict()}

    def load_dict(self, d):
        self.gnn.load_state_dict(d["gnn"])
        self.value_net.load_state_dict(d["value_net"])
        self.action_net.load_state_dict(d["action_net"])
