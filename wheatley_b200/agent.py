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


class PspAgent(AgentHeadMixin, nn.Module):
    """The live agent (psp/models/agent.py:47-226): ``gnn`` = :class:`GnnMP`, ``value_net`` / ``action_net`` =
    ``torch.compile``-wrapped MLPs over ``gnn.features_dim`` (:191-211; their state-dict keys carry ``_orig_mod.``).
    ``preprocess`` is the device batch builder (psp/models/agent.py:298-406 -> wheatley_b200.psp_batch)."""

    def __init__(self, max_n_nodes, max_n_resources, hidden_dim=64, n_layers=6, n_heads=4, n_mlp_layers=3,
                 activation="gelu", graph_pooling="learn", conflicts="clique", net_arch=None, activation_fn="gelu",
                 input_dim=16, residual=True, checkpoint=1, add_rp_edges="frontier_strict", self_loops=True):
        nn.Module.__init__(self)
        from .gnn_mp import Compiled, GnnMP
        net_arch = net_arch or {"vf": [64, 64, 64], "pi": [64, 64, 64]}
        self.conflicts, self.graph_pooling, self.self_loops = conflicts, graph_pooling, self_loops
        self.gnn = GnnMP(input_dim_features_extractor=input_dim, graph_pooling=graph_pooling,
                         max_n_nodes=max_n_nodes, max_n_resources=max_n_resources,
                         n_mlp_layers_features_extractor=n_mlp_layers, n_layers_features_extractor=n_layers,
                         hidden_dim_features_extractor=hidden_dim, activation_features_extractor=activation,
                         n_attention_heads=n_heads, residual=residual, normalize=False, conflicts=conflicts,
                         edge_embedding_flavor="sum", layer_pooling="all", factored_rp=False,
                         add_rp_edges=add_rp_edges, add_self_loops=self_loops, vnode=False, rwpe_k=0, rwpe_h=0,
                         graphobs=True, shared_conv=False, pyg=True, checkpoint=checkpoint)
        self.value_net = Compiled(MLP(len(net_arch["vf"]), self.gnn.features_dim, net_arch["vf"][0], 1, False,
                                      activation_fn))
        self.action_net = Compiled(MLP(len(net_arch["pi"]), self.gnn.features_dim, net_arch["pi"][0], 1, False,
                                       activation_fn))

    def preprocess(self, arena, n_graphs=None, defer=False, **kw):
        from .psp_batch import preprocess
        dev = next(self.parameters()).device
        return preprocess(arena, dev, n_graphs, conflicts=self.conflicts, graph_pooling=self.graph_pooling,
                          self_loops=self.self_loops, defer=defer, **kw)

    def save_dict(self):
        return {"gnn": self.gnn.state_dict(), "value_net": self.value_net.state_dict(),
                "action_net": self.action_net.state_dict()}

    def load_dict(self, d):
        self.gnn.load_state_dict(d["gnn"])
        self.value_net.load_state_dict(d["value_net"])
        self.action_net.load_state_dict(d["action_net"])
