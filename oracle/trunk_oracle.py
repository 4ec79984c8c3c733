"""CPU restatement of the JSSP trunk + agent heads around the attention layer:
jssp/models/gnn_dgl.py:339-431 (embedder, L x (EGATConv -> MLP, residual), concat, pool, pad)
and generic/agent.py:190-231 (value / action heads, masked categorical).
TEST INFRASTRUCTURE - see oracle/__init__.py.  Also the timed CPU baseline of bench.py
(``cpu_baseline`` / ``--impl reference``; DGL itself is not installable here).

Built from the pinned pieces: ``OracleEGATConv`` (pinned modulo DGL primitives, eegat.npz),
``generic/mlp.py`` restated below (pinned by mlp.npz) and ``head_oracle`` (pinned by head.npz).
Parameter names equal the reference's, so one ``state_dict`` drives both this oracle and
``wheatley_b200.GnnDGL``."""
import torch
from torch import nn

from . import gat_oracle, head_oracle

_ACT = {"tanh": nn.Tanh, "relu": nn.LeakyReLU, "elu": nn.ELU, "gelu": nn.GELU, "selu": nn.SELU,
        "silu": nn.SiLU}


class OracleMLP(nn.Module):
    """generic/mlp.py:29-77."""

    def __init__(self, n_layers, input_dim, hidden_dim, output_dim, norm, activation):
        super().__init__()
        self.n_layers, self.norm = n_layers, norm
        self.layers = nn.ModuleList()
        if norm:
            self.norms = nn.ModuleList()
        last = input_dim
        for _ in range(n_layers - 1):
            self.layers.append(nn.Linear(last, hidden_dim))
            if norm:
                self.norms.append(nn.LayerNorm(hidden_dim))
            last = hidden_dim
        self.layers.append(nn.Linear(last, output_dim))
        self.activation_layer = (_ACT[activation] if isinstance(activation, str) else activation)()

    def forward(self, x):
        for i in range(self.n_layers - 1):
            x = self.layers[i](x)
            if self.norm:
                x = self.norms[i](x)
            x = self.activation_layer(x)
        return self.layers[-1](x)


class OracleGnnDGL(nn.Module):
    """gat/gatv2 branch of jssp/models/gnn_dgl.py with residual=True, normalize=False.  Pinned by
    tests/golden/gnn_dgl.npz: the reference's own ``GnnDGL.forward`` (:245-431) executed on the DGL shim
    (tests/golden/make_golden.py::golden_gnn_dgl)."""

    def __init__(self, input_dim, graph_pooling, max_n_nodes, max_n_machines, n_mlp_layers, n_layers,
                 hidden_dim, activation, n_heads, n_edge_types=6, conflicts="clique", checkpoint=False):
        super().__init__()
        self.checkpoint = checkpoint        # recompute each layer in backward (one layer's [E, H*D] tensors live)
        self.graph_pooling, self.max_n_nodes, self.n_layers = graph_pooling, max_n_nodes, n_layers
        self.conflicts, self.max_n_machines = conflicts, max_n_machines
        self.node_embedder = nn.Embedding(1 + max_n_machines, input_dim)
        self.edge_embedder = nn.ModuleList([nn.Embedding(n_edge_types, hidden_dim)])
        self.embedder = OracleMLP(n_mlp_layers, input_dim, hidden_dim, hidden_dim, False, activation)
        self.features_extractors = nn.ModuleList()
        self.mlps = nn.ModuleList()
        for _ in range(n_layers):
            self.features_extractors.append(
                gat_oracle.OracleEGATConv(hidden_dim, hidden_dim, hidden_dim, hidden_dim, n_heads))
            self.mlps.append(OracleMLP(n_mlp_layers, hidden_dim * n_heads, hidden_dim * n_heads,
                                       hidden_dim, False, activation))

    def forward(self, src, dst, etype, feat, batch_size, n_nodes, num_task_nodes):
        """``n_nodes`` = nodes per graph (``num_task_nodes / batch_size``, agent_observation.py:47-51)."""
        features = feat
        node_offset = num_task_nodes
        if self.graph_pooling in ("learn", "learninv"):
            node_offset += batch_size
        if self.graph_pooling == "learn":                                   # gnn_dgl.py:341-344
            features = features.clone()
            features[num_task_nodes: num_task_nodes + batch_size] = self.node_embedder.weight[0]
        if self.conflicts == "node":                                        # :346-351
            M = self.max_n_machines
            features = features.clone()
            features[node_offset: node_offset + M * batch_size] = self.node_embedder(
                torch.arange(M).repeat(batch_size) + 1)
        fl = [features]
        features = self.embedder(features)                                  # :355
        fl.append(features)
        emb = self.edge_embedder[0](etype)                                  # :364
        for l in range(self.n_layers):
            def layer_fn(features, emb, l=l):
                h, _ = self.features_extractors[l](src, dst, features, emb)     # :374-376
                return self.mlps[l](h.flatten(start_dim=-2, end_dim=-1))        # :377
            if self.checkpoint and torch.is_grad_enabled():
                from torch.utils.checkpoint import checkpoint
                features = checkpoint(layer_fn, features, emb, use_reentrant=False)
            else:
                features = layer_fn(features, emb)
            features = features + fl[-1]                                    # :395-396
            fl.append(features)
        features = torch.cat(fl, dim=1)                                     # :401
        node_features = features[:num_task_nodes].reshape(batch_size, n_nodes, -1)      # :402-405
        if self.graph_pooling == "max":                                     # :407-409
            graph_embedding = node_features.max(dim=1)[0]
        elif self.graph_pooling == "avg":                                   # :410-412
            graph_embedding = node_features.mean(dim=1)
        else:                                                               # 'learn' / 'learninv' :413-415
            graph_embedding = features[num_task_nodes: num_task_nodes + batch_size]
        node_features = nn.functional.pad(node_features, (0, 0, 0, self.max_n_nodes - n_nodes))   # :420-425
        return node_features, graph_embedding.reshape(batch_size, 1, -1)

    def forward_cat(self, *a):
        """The reference's return value (:427-431): node features next to the repeated graph embedding."""
        node_features, graph_embedding = self.forward(*a)
        return torch.cat((node_features, graph_embedding.expand(node_features.shape)), dim=2)


class OracleAgent(nn.Module):
    """gnn + value_net + action_net (generic/agent.py:143-147) and the PPO loss terms the update
    evaluates per minibatch (alg/ppo.py:575-705, reduced to policy + entropy + value terms)."""

    def __init__(self, gnn, feat_dim, hidden=64, n_mlp_layers=3, activation="gelu"):
        super().__init__()
        self.gnn = gnn
        self.value_net = OracleMLP(n_mlp_layers, feat_dim, hidden, 1, False, activation)
        self.action_net = OracleMLP(n_mlp_layers, feat_dim, hidden, 1, False, activation)

    def evaluate(self, batch_args, action, masks):
        node_features, graph_embedding = self.gnn(*batch_args)
        value = self.value_net(graph_embedding)
        logits = self.action_net(node_features).squeeze(-1)
        _, logprob, entropy, _ = head_oracle.action_and_logprob(logits, action, masks)
        return logprob, entropy, value

    def loss(self, batch_args, action, masks, advantages, returns, ent_coef=0.05, vf_coef=2.0):
        logprob, entropy, value = self.evaluate(batch_args, action, masks)
        return ppo_like_loss(logprob, entropy, value, advantages, returns, ent_coef, vf_coef)


def ppo_like_loss(logprob, entropy, value, advantages, returns, ent_coef=0.05, vf_coef=2.0):
    """Scalar with the gradient structure of alg/ppo.py:600-705 at the first epoch (ratio = 1):
    ``-(adv * exp(logprob - logprob.detach())).mean() - ent_coef * entropy.mean()
    + vf_coef * 0.5 * ((value - returns)**2).mean()``."""
    ratio = torch.exp(logprob - logprob.detach())
    pg = -(advantages * ratio).mean()
    v = 0.5 * ((value.reshape(-1) - returns) ** 2).mean()
    return pg - ent_coef * entropy.mean() + vf_coef * v
