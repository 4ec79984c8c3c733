"""GnnMP / GnnFlat - the LIVE trunk of the reference (psp/models/gnn_mp.py:38-507, psp/models/gnn_flat.py:7-259;
every JSSP / PSP config runs through it today) on the fused CUDA attention layer.

Same constructor arguments, sub-module / parameter names - including the ``_orig_mod.`` infix ``torch.compile``
puts into the reference's state-dict keys (``gnn._orig_mod.features_extractors.<l>.conv.lin_l.weight`` ...,
SURVEY section 5 "Checkpoint") - and return values, so ``Agent.save`` / ``load_state_dict`` round-trip with
reference checkpoints.  Covered configuration = the only one that runs upstream:
``normalize=False, rwpe_k=0, update_edge_features=False, shared_conv=False, hierarchical=False, tokengt=False,
dual_net=False, layer_pooling="all"``; anything else raises like the reference does (AttributeError on the
attributes it never creates) or NotImplementedError.

Differences, all inside the boundary:
* the batched graph arrives as the 9-tuple of ``wheatley_b200.psp_batch.preprocess`` (built on the device);
* edge features stay factored (``PspEdgeEmbedder`` -> ``AffineEdgeFeats``): ``lin_edge`` projects a 16-row table
  and a 5-row affine map instead of ``[E, hidden]`` (psp/models/edge_embedder.py:130-143);
* the residual add (gnn_flat.py:228-230) rides in the MLP's last GEMM epilogue; the per-graph pad to
  ``max_n_nodes`` (gnn_mp.py:459-470: a Python loop of ``F.pad`` + ``stack``) is one indexed copy."""
import torch
from torch import nn

from . import _lib
from .edge_feats import PspEdgeEmbedder
from .graph_conv import GATv2Conv, GraphConv
from .mlp import MLP
from .plan import GraphPlan
from .readout import readout
from .trunk import LayerSpec, fused_trunk, trunk_applicable, trunk_enabled


class Compiled(nn.Module):
    """Name-compatible stand-in for ``torch.compile(module)``: the wrapped module's parameters appear under
    ``_orig_mod.`` exactly as in the reference's OptimizedModule; calls are forwarded (the hot ops are hand-written
    kernels: there is nothing for a tracing compiler to fuse)."""

    def __init__(self, mod):
        super().__init__()
        self._orig_mod = mod

    def forward(self, *a, **k):
        return self._orig_mod(*a, **k)

    def __getattr__(self, name):
        try:
            return super().__getattr__(name)
        except AttributeError:
            return getattr(super().__getattr__("_orig_mod"), name)


class GnnFlat(nn.Module):
    def __init__(self, input_dim_features_extractor, hidden_dim_features_extractor, n_layers_features_extractor,
                 n_mlp_layers_features_extractor, layer_pooling, n_attention_heads, normalize,
                 activation_features_extractor, residual, rwpe_k, rwpe_h, update_edge_features,
                 update_edge_features_pe, shared_conv, edge_embedding_flavor, max_n_resources, add_rp_edges,
                 factored_rp, pyg, checkpoint):
        super().__init__()
        if normalize or rwpe_k != 0 or update_edge_features or shared_conv or not pyg:
            raise NotImplementedError("only the live GnnFlat configuration is on the hot path (SURVEY F13: the other "
                                      "branches reference attributes the reference never creates)")
        self.hidden_dim = hidden_dim_features_extractor
        self.layer_pooling = layer_pooling
        self.rwpe_k, self.rwpe_h = rwpe_k, rwpe_h
        self.update_edge_features, self.update_edge_features_pe = update_edge_features, update_edge_features_pe
        self.n_layers_features_extractor = n_layers_features_extractor
        self.shared_conv, self.pyg, self.normalize, self.residual = shared_conv, pyg, normalize, residual
        self.input_dim_features_extractor = input_dim_features_extractor
        self.checkpoint = checkpoint
        self.features_extractors = nn.ModuleList()
        self.mlps = nn.ModuleList()
        self.edge_embedders = nn.ModuleList()
        for _ in range(n_layers_features_extractor):                          # gnn_flat.py:96-125
            self.features_extractors.append(GraphConv(self.hidden_dim, self.hidden_dim,
                                                      num_heads=n_attention_heads, edge_scoring=False, pyg=True))
            self.mlps.append(MLP(n_layers=n_mlp_layers_features_extractor,
                                 input_dim=self.hidden_dim * n_attention_heads,
                                 hidden_dim=self.hidden_dim * n_attention_heads, output_dim=self.hidden_dim,
                                 norm=False, activation=activation_features_extractor))
            self.edge_embedders.append(PspEdgeEmbedder(edge_embedding_flavor, max_n_resources, self.hidden_dim,
                                                       add_rp_edges, factored_rp, n_mlp_layers_features_extractor,
                                                       activation_features_extractor))

    def _layer(self, layer, g, features, prev):
        h, _ = self.features_extractors[layer](g._graph, features, self.edge_embedders[layer](g))   # :204-208
        # :209 and the residual add of :228-230, fused into the MLP's last GEMM epilogue
        return self.mlps[layer](h, residual=prev if self.residual else None)

    def _trunk_specs(self, g):
        """Per-layer parameter lists for ``wheatley_b200.trunk``: ``lin_l | lin_r`` (+ biases), ``att``, the output
        ``bias``, and the layer's edge embedder folded through ``lin_edge`` into a 16-row table and a 5-row affine
        map (psp/models/edge_embedder.py:130-143)."""
        specs, ef0 = [], None
        for fe, mlp, emb in zip(self.features_extractors, self.mlps, self.edge_embedders):
            conv = fe.conv
            if (mlp.norm or not isinstance(mlp.activation_layer, nn.GELU)
                    or getattr(mlp.activation_layer, "approximate", "none") != "none"
                    or not isinstance(conv, GATv2Conv) or conv.lin_edge is None or conv.lin_l.bias is None):
                return None, None
            ef = emb(g)
            if ef.attrs is None or ef.attrs.shape[1] > _lib.MAX_ATTR:
                return None, None
            ef0 = ef0 or ef
            W = conv.lin_edge.weight
            specs.append(LayerSpec([conv.lin_l.weight, conv.lin_r.weight], [conv.lin_l.bias, conv.lin_r.bias],
                                   conv.att, conv.bias, nn.functional.linear(ef.weight, W),
                                   (W @ ef.attr_weight).t(), [lin.weight for lin in mlp.layers],
                                   [lin.bias for lin in mlp.layers]))
        return specs, ef0

    def forward(self, g, features, pe):
        plan = getattr(g._graph, "plan", None)
        if self.checkpoint == 1 and isinstance(plan, GraphPlan) and trunk_enabled():
            specs, ef0 = self._trunk_specs(g)
            H = self.features_extractors[0].conv.heads
            if specs is not None and trunk_applicable(features, specs, H, self.hidden_dim):
                # all L layers behind one autograd node / one C-ABI call per pass (wheatley_b200.trunk)
                outs = fused_trunk(plan, plan.sort_edge_ints(ef0.etype), plan.sort_edge_rows(ef0.attrs), features,
                                   specs, heads=H, dim=self.hidden_dim, slope=0.2, eps=1e-16,
                                   residual=self.residual)
                return [features] + outs if self.layer_pooling == "all" else outs[-1]
        features_list = [features]                                            # :155-161
        previous = features
        for layer in range(self.n_layers_features_extractor):
            if layer % self.checkpoint:                                       # :192-202 activation checkpointing
                from torch.utils.checkpoint import checkpoint
                features = checkpoint(self._layer, layer, g, features, previous, use_reentrant=False)
            else:
                features = self._layer(layer, g, features, previous)
            previous = features
            if self.layer_pooling == "all":
                features_list.append(features)                                # :249-253
        return features_list if self.layer_pooling == "all" else features


class GnnMP(nn.Module):
    def __init__(self, input_dim_features_extractor, graph_pooling, max_n_nodes, max_n_resources,
                 n_mlp_layers_features_extractor, n_layers_features_extractor, hidden_dim_features_extractor,
                 activation_features_extractor, n_attention_heads, residual=True, normalize=False, conflicts="att",
                 edge_embedding_flavor="sum", layer_pooling="all", factored_rp=False, add_rp_edges="all",
                 add_self_loops=False, vnode=False, update_edge_features=False, update_edge_features_pe=False,
                 rwpe_k=16, rwpe_h=16, rwpe_cache=None, graphobs=False, shared_conv=True, pyg=True,
                 hierarchical=False, tokengt=False, checkpoint=1, dual_net=False):
        super().__init__()
        if hierarchical or tokengt or dual_net or vnode or graph_pooling == "gap" or normalize:
            raise NotImplementedError("hierarchical / TokenGT / dual-net / vnode / gap / normalised trunks are not "
                                      "the graph-attention hot path (SURVEY section 2)")
        self.pyg, self.conflicts, self.residual, self.normalize = pyg, conflicts, residual, normalize
        self.max_n_nodes = max_n_nodes
        self.input_dim_features_extractor = input_dim_features_extractor
        self.layer_pooling = layer_pooling
        self.rwpe_k = rwpe_k
        self.rwpe_h = 0 if rwpe_k == 0 else rwpe_h
        self.hierarchical, self.tokengt, self.dual_net = hierarchical, tokengt, dual_net
        self.node_type_size = 2
        if layer_pooling == "all":                                            # gnn_mp.py:100-105
            self.features_dim = (input_dim_features_extractor
                                 + (hidden_dim_features_extractor + self.rwpe_h) * (n_layers_features_extractor + 1)
                                 + self.rwpe_h)
        else:
            self.features_dim = hidden_dim_features_extractor + self.rwpe_h
        self.graphobs, self.factored_rp, self.add_rp_edges = graphobs, factored_rp, add_rp_edges
        self.max_n_resources = max_n_resources
        self.add_self_loops = True
        self.hidden_dim = hidden_dim_features_extractor
        self.graph_pooling = graph_pooling
        self.n_layers_features_extractor = n_layers_features_extractor
        self.vnode, self.shared_conv, self.checkpoint = vnode, shared_conv, checkpoint
        self.edge_embedding_flavor = edge_embedding_flavor
        H2 = hidden_dim_features_extractor - self.node_type_size
        self.node_type_embedder = Compiled(nn.Embedding(4, self.node_type_size))          # :135-139
        self.pool_node_embedder = Compiled(nn.Embedding(1, H2))                           # :141-145
        if conflicts == "node":
            self.resource_node_embedder = Compiled(nn.Linear(3, H2))                      # :178-183
        self.gnn = Compiled(GnnFlat(input_dim_features_extractor, hidden_dim_features_extractor,
                                    n_layers_features_extractor, n_mlp_layers_features_extractor, layer_pooling,
                                    n_attention_heads, normalize, activation_features_extractor, residual, rwpe_k,
                                    self.rwpe_h, update_edge_features, update_edge_features_pe, shared_conv,
                                    edge_embedding_flavor, max_n_resources, add_rp_edges, factored_rp, pyg,
                                    checkpoint))                                          # :246-266
        self.features_embedder = Compiled(MLP(n_layers=n_mlp_layers_features_extractor,
                                              input_dim=input_dim_features_extractor, hidden_dim=H2,
                                              output_dim=H2, norm=normalize,
                                              activation=activation_features_extractor))  # :292-300

    def embed_nodes(self, orig_features, num_nodes, poolnodes, resource_nodes, res_cal_id):
        """gnn_mp.py:342-398: the ``[N_total, hidden]`` input of the first layer, assembled with ONE
        concatenation per node class (the reference fills a ``torch.empty`` buffer slice by slice)."""
        dev, dt = orig_features.device, orig_features.dtype
        nte = self.node_type_embedder._orig_mod.weight
        emb = self.features_embedder(orig_features[:num_nodes])                            # :349-351
        parts = [torch.cat((nte[2].expand(num_nodes, -1), emb), dim=1)]                   # :346-351
        n_res = 0 if resource_nodes is None else len(resource_nodes)
        n_pool = 0 if poolnodes is None else len(poolnodes)
        # node order of the builder: tasks | resource nodes | pool nodes (rewiring.py:314-361)
        if self.conflicts == "node" and n_res:                                             # :392-398
            r = self.resource_node_embedder(res_cal_id.to(device=dev, dtype=dt))
            parts.append(torch.cat((nte[1].expand(n_res, -1), r), dim=1))
        if n_pool:                                                                         # :353-359
            p = self.pool_node_embedder._orig_mod.weight[0]
            parts.append(torch.cat((nte[0], p)).expand(n_pool, -1))
        return torch.cat(parts, dim=0) if len(parts) > 1 else parts[0]

    def forward(self, obs):
        (g, batch_size, num_nodes, n_nodes, poolnodes, resource_nodes, vnodes, res_cal_id, pe) = obs
        g = g.to(next(self.parameters()).device)                                           # :331
        features = g.ndata("feat")
        features[:, 2:4] = 0                                                               # :337 (in place, like the reference)
        orig_features = features
        n_total = orig_features.shape[0]
        n_res = 0 if resource_nodes is None else len(resource_nodes)
        n_pool = 0 if poolnodes is None else len(poolnodes)
        if num_nodes + n_res + n_pool != n_total:
            raise _lib.B200GatError("node layout is not tasks | resource nodes | pool nodes")
        if self.graph_pooling in ("learn", "learninv") and n_pool == 0:
            raise _lib.B200GatError("learned pooling without pool nodes")
        features = self.embed_nodes(orig_features, num_nodes, poolnodes, resource_nodes, res_cal_id)
        features = self.gnn(g, features, pe)                                               # :414
        # :419-500 in one pass (b200gat_readout_fwd): the concatenation of all layers is never materialised, every
        # graph's rows land in their padded slot, max / avg pooling (first-maximum gradient like torch.max on the
        # CPU) or the learned pool node's row becomes the graph embedding
        sources = [orig_features] + features if self.layer_pooling == "all" else [features]
        counts = [n_nodes] * batch_size if isinstance(n_nodes, int) else list(n_nodes)
        if self.graph_pooling not in ("learn", "learninv", "max", "avg"):
            raise Exception(f"Graph pooling {self.graph_pooling} not recognized. Only accepted pooling are max and avg")
        node_features, graph_embedding = readout(sources, counts, self.max_n_nodes, self.graph_pooling,
                                                 pool_row0=num_nodes + n_res)
        return node_features, graph_embedding.reshape(batch_size, 1, -1)
