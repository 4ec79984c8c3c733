"""JsspAgent: ``gnn`` + ``value_net`` + ``action_net`` with the reference attribute names
(generic/agent.py:143-147, jssp/models/agent.py:70-121), so ``Agent.save``'s checkpoint layout
``{"gnn": ..., "value_net": ..., "action_net": ...}`` (generic/agent.py:158-172) round-trips.
The three head methods come from :class:`AgentHeadMixin` (generic/agent.py:190-277)."""
import torch
from torch import nn

from . import _lib
from .agent_head import AgentHeadMixin
from .agent_step import STATS, StepSpec, agent_step
from .functional import _mask_u8
from .gnn_dgl import GnnDGL
from .mlp import MLP
from .readout import row_starts
from .trunk import trunk_applicable, trunk_cfg, trunk_enabled


def _gelu_linears(mlp):
    """The ``nn.Linear`` layers of a GELU MLP without LayerNorm (the configuration the fused calls cover), else
    None."""
    mlp = getattr(mlp, "_orig_mod", mlp)
    if mlp.norm or not isinstance(mlp.activation_layer, nn.GELU) or \
            getattr(mlp.activation_layer, "approximate", "none") != "none":
        return None
    return list(mlp.layers)


def _wb(layers):
    out = []
    for lin in layers:
        out += [lin.weight, lin.bias]
    return out


class PpoMinibatchMixin:
    """``ppo_minibatch``: the block alg/ppo.py:575-711 of the reference's update loop as ONE call -
    ``get_action_and_value`` on the minibatch's observations and actions, the clipped-PPO loss and its statistics.
    Returns ``(loss, stats)``: ``loss.backward()`` is the whole backward pass; ``stats`` is a device tensor in the
    order of ``wheatley_b200.agent_step.STATS`` (what the reference reads back with ``.item()`` one value at a time).
    One autograd node / one C-ABI call per pass when the configuration is covered (``wheatley_b200.agent_step``),
    the module-by-module path (same kernels, same numbers) otherwise."""

    def ppo_minibatch(self, obs, actions, action_masks, old_logprobs, advantages, returns, old_values=None, *,
                      clip_coef=0.2, norm_adv=True, clip_vloss=False, critic_loss="l2", ent_coef=0.0, vf_coef=0.5,
                      sharded=False, group=None):
        import torch.distributed as dist
        from .ppo import adv_moments, ppo_loss
        cfg = dict(clip_coef=clip_coef, norm_adv=norm_adv, clip_vloss=clip_vloss, critic_loss=critic_loss,
                   ent_coef=ent_coef, vf_coef=vf_coef)
        built = self._step_inputs(obs) if trunk_enabled() else None
        if built is None:
            _, lp, ent, val, _ = self.get_action_and_value(obs, actions, action_masks)
            return ppo_loss(lp, ent, val, old_logprobs, advantages, returns, old_values, sharded=sharded,
                            group=group, **cfg)
        spec_kw, ext, xin, flat, n_logits = built
        if sharded and dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            cfg["moments"] = adv_moments(advantages)
            dist.all_reduce(cfg["moments"], op=dist.ReduceOp.SUM, group=group)
        spec = StepSpec(ppo=cfg, **spec_kw)
        mask = _mask_u8(action_masks, n_logits, xin.device)
        loss, stats, _aux = agent_step(spec, ext, xin, actions, mask, old_logprobs, advantages, returns, old_values,
                                       flat)
        return loss, stats


class JsspAgent(PpoMinibatchMixin, AgentHeadMixin, nn.Module):
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

    def _step_inputs(self, b):
        """Operands of ``wheatley_b200.agent_step`` for a :class:`JsspBatch`, or None when not covered."""
        gnn = self.gnn
        emb, act, val = _gelu_linears(gnn.embedder), _gelu_linears(self.action_net), _gelu_linears(self.value_net)
        if emb is None or act is None or val is None or gnn.graph_pooling not in ("learn", "learninv", "max", "avg"):
            return None
        if b.num_task_nodes != b.batch_size * b.n_nodes:
            return None
        x_in = gnn.input_features(b)
        efeats = gnn.edge_features(b)
        specs = gnn._trunk_specs(efeats)
        H, D = gnn.features_extractors[0]._num_heads, gnn.hidden_dim
        probe = x_in.new_empty((x_in.shape[0], D))
        if specs is None or not trunk_applicable(probe, specs, H, D) or any(lin.weight.shape[0] % 4 for lin in emb[:-1]):
            return None
        plan = b.plan
        if plan.stats_host()[0] > 0:
            return None                      # the module path raises the reference's DGLError
        etype_sorted = plan.sort_edge_ints(b.etype)
        if getattr(b, "row_start", None) is None:
            b.row_start = row_starts([b.n_nodes] * b.batch_size, x_in.device)
        L = len(specs)
        flat = _wb(emb)
        for sp in specs:
            flat += sp.flat()
        flat += _wb(act) + _wb(val)
        kw = dict(view=plan.clique_view(etype_sorted), etype_sorted=etype_sorted, attr_sorted=None,
                  trunk_cfg=trunk_cfg(specs, H, D, 0.01, 0.0, gnn.residual), n_embedder=len(emb),
                  n_action=len(act), n_value=len(val),
                  src_kinds=[_lib.SRC_EXTERNAL, _lib.SRC_X0] + [_lib.SRC_LAYER0 + l for l in range(L)],
                  src_widths=[x_in.shape[1]] + [D] * (L + 1), counts=[b.n_nodes] * b.batch_size,
                  row_start=b.row_start, max_n_nodes=gnn.max_n_nodes, pooling=gnn.graph_pooling,
                  pool_row0=b.num_task_nodes)
        return kw, x_in, x_in, flat, gnn.max_n_nodes

    def save_dict(self):
        """generic/agent.py:158-172 layout (the two pickled specs are the caller's)."""
        return {"gnn": self.gnn.state_dict(), "value_net": self.value_net.state_dict(),
                "action_net": self.action_net.state_dict()}

    def load_dict(self, d):
        self.gnn.load_state_dict(d["gnn"])
        self.value_net.load_state_dict(d["value_net"])
        self.action_net.load_state_dict(d["action_net"])


class PspAgent(PpoMinibatchMixin, AgentHeadMixin, nn.Module):
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

    def _step_inputs(self, obs):
        """Operands of ``wheatley_b200.agent_step`` for the 9-tuple of ``preprocess``, or None when not covered."""
        from .plan import GraphPlan
        (g, batch_size, num_nodes, n_nodes, poolnodes, resource_nodes, _vn, res_cal_id, pe) = obs
        gnn = self.gnn
        flatnet = gnn.gnn._orig_mod
        act, val = _gelu_linears(self.action_net), _gelu_linears(self.value_net)
        plan = getattr(g._graph, "plan", None)
        if (act is None or val is None or not isinstance(plan, GraphPlan) or flatnet.checkpoint != 1
                or gnn.graph_pooling not in ("learn", "learninv", "max", "avg")):
            return None
        specs, ef0 = flatnet._trunk_specs(g)
        H, D = flatnet.features_extractors[0].conv.heads, gnn.hidden_dim
        features = g.ndata("feat")
        features[:, 2:4] = 0                                                   # gnn_mp.py:337
        n_res = 0 if resource_nodes is None else len(resource_nodes)
        n_pool = 0 if poolnodes is None else len(poolnodes)
        if num_nodes + n_res + n_pool != features.shape[0]:
            return None
        if gnn.graph_pooling in ("learn", "learninv") and n_pool == 0:
            return None
        x0 = gnn.embed_nodes(features, num_nodes, poolnodes, resource_nodes, res_cal_id)
        if specs is None or not trunk_applicable(x0, specs, H, D):
            return None
        counts = [n_nodes] * batch_size if isinstance(n_nodes, int) else list(n_nodes)
        L = len(specs)
        all_layers = gnn.layer_pooling == "all"
        flat = []
        for sp in specs:
            flat += sp.flat()
        flat += _wb(act) + _wb(val)
        kinds = ([_lib.SRC_EXTERNAL, _lib.SRC_X0] + [_lib.SRC_LAYER0 + l for l in range(L)] if all_layers
                 else [_lib.SRC_LAYER0 + L - 1])
        widths = ([features.shape[1]] + [D] * (L + 1)) if all_layers else [D]
        kw = dict(view=plan, etype_sorted=plan.sort_edge_ints(ef0.etype), attr_sorted=plan.sort_edge_rows(ef0.attrs),
                  trunk_cfg=trunk_cfg(specs, H, D, 0.2, 1e-16, flatnet.residual), n_embedder=0, n_action=len(act),
                  n_value=len(val), src_kinds=kinds, src_widths=widths, counts=counts,
                  row_start=row_starts(counts, x0.device), max_n_nodes=gnn.max_n_nodes, pooling=gnn.graph_pooling,
                  pool_row0=num_nodes + n_res)
        return kw, (features if all_layers else None), x0, flat, gnn.max_n_nodes

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
