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


# --------------------------------------------------------------------------------------------------------------
# PSP / PyG flavour: psp/models/gnn_mp.py GnnMP + psp/models/gnn_flat.py GnnFlat (the live trunk, SURVEY F13:
# normalize_gnn=False, rwpe_k=0, update_edge_features=False, shared_conv=False, checkpoint=1, layer_pooling="all").
# Pinned by tests/golden/gnn_mp.npz (the reference's GnnMP.forward executed, make_golden.py::golden_gnn_mp).
class _OrigMod(nn.Module):
    """Name-compatible stand-in for ``torch.compile``'s OptimizedModule: parameters live under ``_orig_mod.``
    (psp/models/gnn_mp.py:139,145,183,266,300), the call is forwarded."""

    def __init__(self, mod):
        super().__init__()
        self._orig_mod = mod

    def forward(self, *a, **k):
        return self._orig_mod(*a, **k)


class OraclePspEdgeEmbedder(nn.Module):
    """psp/models/edge_embedder.py:25-143, ``edge_embedding_flavor == "sum"``, ``add_rp_edges != "none"``,
    ``factored_rp=False`` (pinned by edge_embedder.npz)."""

    def __init__(self, hidden_dim):
        super().__init__()
        self.edge_type_embedder = nn.Embedding(16, hidden_dim)
        self.rc_att_embedder = nn.Linear(2, hidden_dim)
        self.rp_att_embedder = nn.Linear(3, hidden_dim)

    def forward(self, etype, att_rc, att_rp):                               # :130-143
        return self.edge_type_embedder(etype) + self.rc_att_embedder(att_rc) + self.rp_att_embedder(att_rp)


class _OracleGraphConv(nn.Module):
    """psp/graph/graph_conv.py:8-60 with pyg=True: ``self.conv`` = GATv2Conv(in, out, heads, edge_dim=in)."""

    def __init__(self, in_dim, out_dim, heads):
        super().__init__()
        self.conv = gat_oracle.OracleGATv2Conv(in_dim, out_dim, heads, edge_dim=in_dim)

    def forward(self, edge_index, x, edge_attr):
        return self.conv(x, edge_index, edge_attr)


class OracleGnnFlat(nn.Module):
    """psp/models/gnn_flat.py:151-259 on the live configuration."""

    def __init__(self, hidden_dim, n_layers, n_mlp_layers, n_heads, activation, residual=True):
        super().__init__()
        self.hidden_dim, self.n_layers, self.residual = hidden_dim, n_layers, residual
        self.features_extractors = nn.ModuleList()
        self.mlps = nn.ModuleList()
        self.edge_embedders = nn.ModuleList()
        for _ in range(n_layers):                                           # :96-125
            self.features_extractors.append(_OracleGraphConv(hidden_dim, hidden_dim, n_heads))
            self.mlps.append(OracleMLP(n_mlp_layers, hidden_dim * n_heads, hidden_dim * n_heads, hidden_dim,
                                       False, activation))
            self.edge_embedders.append(OraclePspEdgeEmbedder(hidden_dim))

    def forward(self, edge_index, etype, att_rc, att_rp, features):
        fl = [features]                                                     # :155-161
        for l in range(self.n_layers):
            ef = self.edge_embedders[l](etype, att_rc, att_rp)              # :207
            features = self.features_extractors[l](edge_index, features, ef)   # :204-208
            features = self.mlps[l](features)                               # :209
            if self.residual:
                features = features + fl[-1][:, : self.hidden_dim]          # :228-230
            fl.append(features)                                             # :249-253
        return fl


class OracleGnnMP(nn.Module):
    """psp/models/gnn_mp.py:318-507: node-type / pool / resource embedders, GnnFlat, concat of all layers,
    per-graph pooling, pad to ``max_n_nodes``; returns ``(node_features [B, max_n_nodes, F], graph_embedding
    [B, 1, F])``."""

    def __init__(self, input_dim, graph_pooling, max_n_nodes, n_mlp_layers, n_layers, hidden_dim, activation,
                 n_heads, conflicts="clique", residual=True):
        super().__init__()
        self.graph_pooling, self.max_n_nodes, self.conflicts = graph_pooling, max_n_nodes, conflicts
        self.hidden_dim, self.node_type_size = hidden_dim, 2
        self.node_type_embedder = _OrigMod(nn.Embedding(4, 2))              # :135-139
        self.pool_node_embedder = _OrigMod(nn.Embedding(1, hidden_dim - 2))  # :141-145
        if conflicts == "node":
            self.resource_node_embedder = _OrigMod(nn.Linear(3, hidden_dim - 2))   # :178-183
        self.gnn = _OrigMod(OracleGnnFlat(hidden_dim, n_layers, n_mlp_layers, n_heads, activation, residual))
        self.features_embedder = _OrigMod(OracleMLP(n_mlp_layers, input_dim, hidden_dim - 2, hidden_dim - 2,
                                                    False, activation))     # :292-300

    def forward(self, edge_index, etype, att_rc, att_rp, feat, batch_size, num_nodes, n_nodes, poolnodes,
                resource_nodes, res_cal_id):
        """``n_nodes``: list of per-graph node counts (an int when ``batch_size == 1``)."""
        orig = feat.clone()
        orig[:, 2:4] = 0                                                    # :337
        ts = self.node_type_size
        features = torch.zeros((orig.shape[0], self.hidden_dim), dtype=orig.dtype)
        nte = self.node_type_embedder._orig_mod.weight
        features[:num_nodes, :ts] = nte[2]                                  # :346-348
        features[:num_nodes, ts:] = self.features_embedder(orig[:num_nodes])   # :349-351
        if self.graph_pooling in ("learn", "learninv") and poolnodes is not None and len(poolnodes):
            # :353-359 (the reference fills these rows for "learn" only and leaves them uninitialised for
            # "learninv")
            features[poolnodes, :ts] = nte[0]
            features[poolnodes, ts:] = self.pool_node_embedder._orig_mod.weight[0]
        if self.conflicts == "node":                                        # :392-398
            features[resource_nodes, :ts] = nte[1]
            features[resource_nodes, ts:] = self.resource_node_embedder(res_cal_id)
        fl = self.gnn(edge_index, etype, att_rc, att_rp, features)          # :414
        features = torch.cat([orig] + fl, dim=-1)                           # :419-425
        node_features = features[:num_nodes]
        counts = [n_nodes] * batch_size if isinstance(n_nodes, int) else list(n_nodes)
        chunks = torch.split(node_features, counts)
        if self.graph_pooling == "max":                                     # :430-440, :475-477
            graph_embedding = torch.stack([c.max(dim=0)[0] for c in chunks])
        elif self.graph_pooling == "avg":                                   # :441-452, :478-482
            graph_embedding = torch.stack([c.mean(dim=0) for c in chunks])
        else:                                                               # :453-457, :485-489
            graph_embedding = features[poolnodes]
        node_features = torch.stack([nn.functional.pad(c, (0, 0, 0, self.max_n_nodes - c.shape[0]))
                                     for c in chunks])                      # :459-470, :495-500
        return node_features, graph_embedding.reshape(batch_size, 1, -1)
