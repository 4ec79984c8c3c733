"""GnnDGL - the JSSP trunk of jssp/models/gnn_dgl.py:36-431 (``gconv_type in {"gat","gatv2"}``
branch, the default) on the fused CUDA attention layer.

Same constructor arguments, sub-module / parameter names (``node_embedder``,
``edge_embedder.0``, ``embedder.layers.*``, ``features_extractors.<l>.fc_*``, ``mlps.<l>.layers.*``)
and return layout, so ``Agent.save`` / ``load_state_dict`` round-trip with reference
checkpoints.  Differences, all inside the boundary:

* the batched graph arrives as a :class:`JsspBatch` (built once per batch on the device by
  ``wheatley_b200.batch``) instead of being rebuilt with per-graph Python loops on the CPU
  (``agent_observation.py:191-225``, ``gnn_dgl.py:272-286``);
* edge features stay factored: ``EGATConv`` gets ``TypedEdgeFeats(type, edge_embedder[0].weight)``
  so ``fc_fij`` projects the T-row table, not ``[E, hidden]`` (:364, generic/eegatconv.py:262);
* the ``[E, H, D]`` edge output of the layer is never materialised (its only caller drops it, :374-379).
"""
import torch
from torch import nn

from .edge_feats import TypedEdgeFeats, edge_tables
from .eegatconv import DGLError, EGATConv
from ._lib import B200GatError
from .mlp import MLP
from .readout import readout, row_starts
from .trunk import LayerSpec, fused_trunk, trunk_applicable


class _SplitRows(torch.autograd.Function):
    """``(x[:n], x[n:])`` as views; backward = one ``cat`` of the two gradients."""

    @staticmethod
    def forward(ctx, x, n):
        ctx.n, ctx.shape = n, x.shape
        ctx.set_materialize_grads(False)
        return x[:n], x[n:]

    @staticmethod
    def backward(ctx, ga, gb):
        if ga is None and gb is None:
            return None, None
        n, shape = ctx.n, ctx.shape
        ref = ga if ga is not None else gb
        if ga is None:
            ga = ref.new_zeros((n,) + tuple(shape[1:]))
        if gb is None:
            gb = ref.new_zeros((shape[0] - n,) + tuple(shape[1:]))
        return torch.cat((ga, gb), 0), None


class JsspBatch:
    """What ``GnnDGL.forward`` needs from one batched observation."""

    def __init__(self, plan, etype, feat, batch_size, n_nodes, num_task_nodes, poolnodes=None,
                 machinenodes=None, prepare=True):
        self.plan, self.etype, self.feat = plan, etype, feat
        self.batch_size, self.n_nodes, self.num_task_nodes = batch_size, n_nodes, num_task_nodes
        self.poolnodes, self.machinenodes = poolnodes, machinenodes
        if prepare and hasattr(plan, "prepare"):
            plan.prepare(etype)         # clique view + host counts with one device->host copy, before the trunk


class GnnDGL(nn.Module):
    def __init__(self, input_dim_features_extractor, gconv_type, graph_pooling, graph_has_relu,
                 max_n_nodes, max_n_machines, n_mlp_layers_features_extractor,
                 n_layers_features_extractor, hidden_dim_features_extractor,
                 activation_features_extractor, n_attention_heads, residual=True, normalize=False,
                 conflicts="att", no_tct=False, mid_in_edges=False, split_output=False):
        super().__init__()
        # split_output=True returns the (node_features, graph_embedding) pair the generic Agent
        # unpacks (generic/agent.py:193); False keeps the reference's concatenated tensor (:427-431)
        self.split_output = split_output
        if gconv_type not in ("gat", "gatv2"):
            raise NotImplementedError("only the gat/gatv2 trunk is on the hot path (SURVEY 8a)")
        if normalize:
            raise NotImplementedError("normalize=True is not on the live path (args.py:446-448)")
        self.conflicts, self.residual, self.normalize = conflicts, residual, normalize
        self.max_n_nodes, self.max_n_machines = max_n_nodes, max_n_machines
        self.features_dim = 2 * (input_dim_features_extractor
                                 + hidden_dim_features_extractor * (n_layers_features_extractor + 1))
        self.hidden_dim = hidden_dim_features_extractor
        self.gconv_type, self.graph_has_relu, self.graph_pooling = gconv_type, graph_has_relu, graph_pooling
        self.n_layers_features_extractor = n_layers_features_extractor
        self.no_tct, self.mid_in_edges = no_tct, mid_in_edges
        self.features_extractors = nn.ModuleList()
        self.edge_embedder = nn.ModuleList()
        self.node_embedder = nn.Embedding(1 + max_n_machines, input_dim_features_extractor)
        if conflicts in ("clique", "node"):                      # gnn_dgl.py:89-113
            nmachineid = max_n_machines if conflicts == "clique" else 2 * max_n_machines
            n_types = 5 + nmachineid + 1 if mid_in_edges else 5 + 1
        else:
            n_types = 5
        self.edge_embedder.append(nn.Embedding(n_types, hidden_dim_features_extractor))
        self.embedder = MLP(n_mlp_layers_features_extractor, input_dim_features_extractor,
                            hidden_dim_features_extractor, hidden_dim_features_extractor,
                            norm=normalize, activation=activation_features_extractor)
        self.mlps = nn.ModuleList()
        for _ in range(n_layers_features_extractor):
            self.features_extractors.append(EGATConv(self.hidden_dim, self.hidden_dim, self.hidden_dim,
                                                     self.hidden_dim, num_heads=n_attention_heads))
            self.mlps.append(MLP(n_mlp_layers_features_extractor,
                                 hidden_dim_features_extractor * n_attention_heads,
                                 hidden_dim_features_extractor * n_attention_heads,
                                 hidden_dim_features_extractor, norm=normalize,
                                 activation=activation_features_extractor))

    def node_embeddings(self, b: JsspBatch):
        """Concatenation of all layers' node embeddings ``[N_total, F]`` (gnn_dgl.py:339-401)."""
        return torch.cat(self.layer_embeddings(b), dim=1)                       # :401

    def input_features(self, b: JsspBatch):
        """Raw node features with the pool-node / machine-node rows overwritten by their embeddings (:341-351)."""
        features = b.feat
        if self.graph_pooling == "learn" and b.poolnodes is not None:           # :341-344
            features = features.clone()
            features[b.poolnodes] = self.node_embedder.weight[0].to(features.dtype)
        if self.conflicts == "node" and b.machinenodes is not None:             # :346-351
            ids = (torch.arange(self.max_n_machines, device=features.device) + 1).repeat(b.batch_size)
            features = features.clone()
            features[b.machinenodes] = self.node_embedder(ids)
        return features

    def edge_features(self, b: JsspBatch):
        """``edge_embedder[0](g.edata["type"])`` (:364) kept factored, with every layer's projected edge-type table
        from one GEMM (each layer would do its own 6-row one)."""
        efeats = TypedEdgeFeats(b.etype, self.edge_embedder[0].weight)
        efeats.tables = edge_tables(self.edge_embedder[0].weight,
                                    [fe.fc_fij.weight for fe in self.features_extractors],
                                    [fe.bias for fe in self.features_extractors])
        return efeats

    def layer_embeddings(self, b: JsspBatch):
        """The list the reference concatenates at :401: input features, embedder output, every layer's output."""
        features = self.input_features(b)
        features_list = [features]
        features = self.embedder(features)                                      # :355
        features_list.append(features)
        efeats = self.edge_features(b)
        specs = self._trunk_specs(efeats)
        if specs is not None and trunk_applicable(features, specs, self.features_extractors[0]._num_heads, self.hidden_dim):
            # all L layers behind one autograd node / one C-ABI call per pass (wheatley_b200.trunk)
            plan = b.plan
            if plan.stats_host()[0] > 0:                                        # generic/eegatconv.py:238-247
                raise DGLError("There are 0-in-degree nodes in the graph, output for those nodes will be invalid. "
                               "Adding self-loop on the input graph by calling `g = dgl.add_self_loop(g)` will "
                               "resolve the issue.")
            etype_sorted = plan.sort_edge_ints(b.etype)
            outs = fused_trunk(plan.clique_view(etype_sorted), etype_sorted, None, features, specs,
                               heads=self.features_extractors[0]._num_heads, dim=self.hidden_dim, slope=0.01, eps=0.0,
                               residual=self.residual)
            return features_list + outs
        for layer in range(self.n_layers_features_extractor):
            h, _ = self.features_extractors[layer](b.plan, features, efeats)    # :374-376
            # :377 and the residual add of :395-396 (fused into the MLP's last GEMM epilogue)
            features = self.mlps[layer](h.flatten(start_dim=-2, end_dim=-1),
                                        residual=features_list[-1] if self.residual else None)
            features_list.append(features)
        return features_list

    def _trunk_specs(self, efeats):
        """Per-layer parameter lists for ``wheatley_b200.trunk`` (None when a layer is not the GELU / no-LayerNorm
        configuration or the projected edge tables are missing)."""
        if efeats.tables is None:
            return None
        specs = []
        for fe, mlp in zip(self.features_extractors, self.mlps):
            if (mlp.norm or not isinstance(mlp.activation_layer, nn.GELU)
                    or getattr(mlp.activation_layer, "approximate", "none") != "none" or fe._edge_scoring):
                return None
            specs.append(LayerSpec([fe.fc_ni.weight, fe.fc_nj.weight, fe.fc_node_src.weight], None, fe.attn, None,
                                   efeats.tables[id(fe.fc_fij.weight)], None,
                                   [lin.weight for lin in mlp.layers], [lin.bias for lin in mlp.layers]))
        return specs

    def forward_split(self, b: JsspBatch):
        """``(node_features [B, max_n_nodes, F], graph_embedding [B, 1, F])`` - the pair the generic
        ``Agent`` unpacks (generic/agent.py:193)."""
        # :401-425 in one pass (b200gat_readout_fwd): no [N, F] concatenation, no row split, no pad
        # copy - the per-layer tensors are read where they lie and every graph's rows land in their padded slot
        if self.graph_pooling not in ("learn", "learninv", "max", "avg"):
            raise Exception(f"Graph pooling {self.graph_pooling} not recognized.")
        if b.num_task_nodes != b.batch_size * b.n_nodes:
            raise B200GatError("GnnDGL needs equal-sized graphs (the reference reshapes to [B, n_nodes, F], :405)")
        if getattr(b, "row_start", None) is None:        # once per batch, shared by every epoch over it
            b.row_start = row_starts([b.n_nodes] * b.batch_size, b.feat.device)
        node_features, graph_embedding = readout(self.layer_embeddings(b), [b.n_nodes] * b.batch_size,
                                                 self.max_n_nodes, self.graph_pooling, pool_row0=b.num_task_nodes,
                                                 row_start=b.row_start)
        return node_features, graph_embedding.reshape(b.batch_size, 1, -1)

    def forward(self, b: JsspBatch):
        """Reference return layout (:427-431): ``cat(node_features, repeated graph embedding)``."""
        node_features, graph_embedding = self.forward_split(b)
        if self.split_output:
            return node_features, graph_embedding
        return torch.cat((node_features, graph_embedding.expand(node_features.shape)), dim=2)
