"""Generate the golden fixtures under tests/golden/ by EXECUTING THE REFERENCE.

Run in the build container only (it needs /root/reference, which does not exist
on the GPU box):

    python tests/golden/make_golden.py

What is executed verbatim from the reference tree:
  * generic/eegatconv.py  EEGATConv.forward   (on the DGL primitive shims, oracle/shims/dgl)
  * generic/gatv3_conv.py GATv3Conv.forward   (on the PyG primitive shims, oracle/shims/torch_geometric)
  * generic/agent.py      Agent.get_action_and_value / get_action_probs_and_value / predict
  * generic/mlp.py        MLP.forward
  * psp/models/edge_embedder.py PspEdgeEmbedder.forward
  * jssp/utils/utils.py   compute_conflicts_cliques
  * jssp/models/agent_observation.py AgentObservation.from_gym_observation (DGL shim)
  * jssp/models/gnn_dgl.py GnnDGL.forward (DGL shim, dgl.nn.EGATConv := generic/eegatconv.py EEGATConv)
  * psp/graph/graph_conv.py GraphConv.forward / forward_nog (PyG shim; GATv2Conv itself is a restatement)
  * psp/models/agent.py Agent._obs_as_tensor_graph / Agent.preprocess + psp/models/rewiring.py +
    psp/models/agent_graph_observation.py (on oracle/shims/psp_graph.py, a stand-in for the HeteroData container)
  * psp/models/gnn_mp.py GnnMP.forward + psp/models/gnn_flat.py GnnFlat.forward (torch.compile on the eager backend)
  * alg/ppo.py:276-299 (GAE loop) and :582-705 (minibatch loss): the source lines themselves, exec'ed

Outputs: small .npz files (float64 for the layers so that the comparison with
the oracle is tight; float32 where the reference module is float32-only).
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle", "shims"))
sys.path.insert(0, "/root/reference")


def np_(t):
    return t.detach().cpu().numpy()


def rand_graph(gen, n, e, self_loops=True):
    src = torch.randint(0, n, (e,), generator=gen)
    dst = torch.randint(0, n, (e,), generator=gen)
    if self_loops:
        loop = torch.arange(n)
        src = torch.cat([src, loop])
        dst = torch.cat([dst, loop])
    return src, dst


def grads_of(loss, tensors):
    gs = torch.autograd.grad(loss, tensors, allow_unused=True)
    return [torch.zeros_like(t) if g is None else g for g, t in zip(gs, tensors)]


# ---------------------------------------------------------------- a1: EEGAT
def golden_eegat(edge_scoring: bool, fname: str, seed: int):
    import dgl
    from generic.eegatconv import EEGATConv

    gen = torch.Generator().manual_seed(seed)
    N, E0, Fin, Fe, Do, De, H = 13, 40, 6, 7, 5, 5, 3
    src, dst = rand_graph(gen, N, E0)
    E = src.numel()
    torch.manual_seed(seed)
    layer = EEGATConv(Fin, Fe, Do, De, H, bias=True, edge_scoring=edge_scoring).double()
    with torch.no_grad():
        layer.bias.copy_(torch.randn(layer.bias.shape, generator=gen, dtype=torch.float64) * 0.1)
    x = torch.randn(N, Fin, generator=gen, dtype=torch.float64, requires_grad=True)
    ef = torch.randn(E, Fe, generator=gen, dtype=torch.float64, requires_grad=True)
    efe = torch.randn(E, Fe, generator=gen, dtype=torch.float64, requires_grad=True)
    g = dgl.graph((src, dst), num_nodes=N)
    if edge_scoring:
        h, f, e_out, a = layer(g, x, ef, efe, get_attention=True)
    else:
        h, f, a = layer(g, x, ef, get_attention=True)
    R = torch.randn(h.shape, generator=gen, dtype=torch.float64)
    loss = (h * R).sum()
    names = [k for k, _ in layer.named_parameters()]
    params = [p for _, p in layer.named_parameters()]
    gr = grads_of(loss, params + [x, ef])
    out = dict(src=np_(src), dst=np_(dst), x=np_(x), ef=np_(ef), efe=np_(efe), R=np_(R),
               h=np_(h), f=np_(f), a=np_(a), gx=np_(gr[-2]), gef=np_(gr[-1]),
               heads=np.int64(H))
    if edge_scoring:
        out["e_out"] = np_(e_out)
    for k, p, gp in zip(names, params, gr):
        out["p." + k] = np_(p)
        out["g." + k] = np_(gp)
    np.savez_compressed(os.path.join(HERE, fname), **out)
    # the reference raises on zero-in-degree graphs
    g_bad = dgl.graph((torch.tensor([0, 1]), torch.tensor([1, 1])), num_nodes=3)
    try:
        layer(g_bad, x[:3], ef[:2])
        raise SystemExit("expected DGLError")
    except dgl.base.DGLError:
        pass


# ---------------------------------------------------------------- a3: GATv3
def golden_gatv3(fname: str, seed: int):
    from generic.gatv3_conv import GATv3Conv

    gen = torch.Generator().manual_seed(seed)
    N, E0, Fin, C, H = 11, 30, 6, 8, 2
    src, dst = rand_graph(gen, N, E0, self_loops=False)
    src[3], dst[3] = 4, 4                      # one pre-existing self loop (gets removed/re-added)
    ei = torch.stack([src, dst])
    torch.manual_seed(seed)
    layer = GATv3Conv(Fin, C, H, dropout=0, edge_dim=Fin, bias=True).double()
    with torch.no_grad():
        layer.lin_l.bias.copy_(torch.randn(H * C, generator=gen, dtype=torch.float64) * 0.1)
        layer.lin_r.bias.copy_(torch.randn(H * C, generator=gen, dtype=torch.float64) * 0.1)
    assert layer.bias is None                  # reference quirk, generic/gatv3_conv.py:108-113
    x = torch.randn(N, Fin, generator=gen, dtype=torch.float64, requires_grad=True)
    ea = torch.randn(E0, Fin, generator=gen, dtype=torch.float64)
    out, (ei2, alpha) = layer(x, ei, ea, return_attention_weights=True)
    R = torch.randn(out.shape, generator=gen, dtype=torch.float64)
    names = [k for k, _ in layer.named_parameters()]
    params = [p for _, p in layer.named_parameters()]
    gr = grads_of((out * R).sum(), params + [x])
    d = dict(edge_index=np_(ei), x=np_(x), ea=np_(ea), R=np_(R), out=np_(out),
             edge_index_out=np_(ei2), alpha=np_(alpha), gx=np_(gr[-1]), heads=np.int64(H))
    for k, p, gp in zip(names, params, gr):
        d["p." + k] = np_(p)
        d["g." + k] = np_(gp)
    np.savez_compressed(os.path.join(HERE, fname), **d)


# ---------------------------------------------------------------- a6/a7: head
def golden_head(fname: str, seed: int):
    from generic.agent import Agent

    gen = torch.Generator().manual_seed(seed)
    B, NN, F = 6, 17, 12
    spec = types.SimpleNamespace(reward_weights=[1.0], two_hot=None, hl_gauss=None, symlog=False)
    agent = Agent(env_specification=object(), agent_specification=spec)
    node_features = torch.randn(B, NN, F, generator=gen)
    graph_embedding = torch.randn(B, 1, F, generator=gen)
    agent.gnn = lambda x: (node_features, graph_embedding)
    torch.manual_seed(seed)
    agent.value_net = torch.nn.Linear(F, 1)
    agent.action_net = torch.nn.Linear(F, 1)
    masks = torch.rand(B, NN + 5, generator=gen) < 0.4       # wider than logits: cropped (:202-203)
    masks[:, 0] = True
    masks[2, :] = False
    masks[2, 7] = True                                          # single legal action
    action_in = torch.tensor([int(torch.nonzero(masks[b, :NN])[0]) for b in range(B)])
    d = dict(node_features=np_(node_features), graph_embedding=np_(graph_embedding),
             masks=np_(masks), action_in=np_(action_in),
             w_action=np_(agent.action_net.weight), b_action=np_(agent.action_net.bias),
             w_value=np_(agent.value_net.weight), b_value=np_(agent.value_net.bias))
    # deterministic, masked
    a, lp, ent, val, um = agent.get_action_and_value(None, None, masks, deterministic=True)
    d.update(det_action=np_(a), det_logprob=np_(lp), det_entropy=np_(ent), value=np_(val),
             unmasked_logits=np_(um.logits))
    # evaluate given actions (the PPO update call, alg/ppo.py:575-581) + grads wrt logits
    nf = node_features.clone().requires_grad_(True)
    agent.gnn = lambda x: (nf, graph_embedding)
    a2, lp2, ent2, _, _ = agent.get_action_and_value(None, action_in, masks)
    wl = torch.randn(B, generator=gen)
    we = torch.randn(B, generator=gen)
    (g_nf,) = torch.autograd.grad((lp2 * wl).sum() + (ent2 * we).sum(), [nf])
    d.update(eval_logprob=np_(lp2), eval_entropy=np_(ent2), wl=np_(wl), we=np_(we), g_nf=np_(g_nf))
    agent.gnn = lambda x: (node_features, graph_embedding)
    # no mask
    a3, lp3, ent3, _, _ = agent.get_action_and_value(None, action_in, None)
    d.update(nomask_logprob=np_(lp3), nomask_entropy=np_(ent3))
    # sampling: seeded
    torch.manual_seed(1234)
    a4, lp4, _, _, _ = agent.get_action_and_value(None, None, masks, deterministic=False)
    d.update(sample_action=np_(a4), sample_logprob=np_(lp4))
    # pretrain probs (:233-254) and predict (:256-277)
    probs, _ = agent.get_action_probs_and_value(None, masks)
    probs_nm, _ = agent.get_action_probs_and_value(None, None)
    d.update(probs=np_(probs), probs_nomask=np_(probs_nm))
    d["predict_det"] = np_(agent.predict(None, True, masks))
    torch.manual_seed(4321)
    d["predict_sample"] = np_(agent.predict(None, False, masks))
    np.savez_compressed(os.path.join(HERE, fname), **d)


# ---------------------------------------------------------------- a8: MLP
def golden_mlp(fname: str, seed: int):
    from generic.mlp import MLP

    gen = torch.Generator().manual_seed(seed)
    d = {}
    for i, (act, norm) in enumerate([("gelu", False), ("tanh", True), ("relu", False),
                                     ("elu", False), ("selu", False), ("silu", False)]):
        torch.manual_seed(seed + i)
        m = MLP(3, 10, 12, 5, norm, act)
        x = torch.randn(9, 10, generator=gen, requires_grad=True)
        y = m(x)
        R = torch.randn(y.shape, generator=gen)
        params = list(m.parameters())
        gr = grads_of((y * R).sum(), params + [x])
        d[f"{i}.act"] = np.array(act)
        d[f"{i}.norm"] = np.array(norm)
        d[f"{i}.x"], d[f"{i}.y"], d[f"{i}.R"], d[f"{i}.gx"] = np_(x), np_(y), np_(R), np_(gr[-1])
        for (k, p), gp in zip(m.named_parameters(), gr):
            d[f"{i}.p.{k}"] = np_(p)
            d[f"{i}.g.{k}"] = np_(gp)
    np.savez_compressed(os.path.join(HERE, fname), **d)


# ---------------------------------------------------------------- a9: edge embedder
def golden_edge_embedder(fname: str, seed: int):
    from psp.models.edge_embedder import PspEdgeEmbedder

    gen = torch.Generator().manual_seed(seed)
    E, Hd = 23, 8
    torch.manual_seed(seed)
    emb = PspEdgeEmbedder("sum", 4, Hd, "all", False, 3, "gelu")
    etype = torch.randint(0, 16, (E,), generator=gen)
    att_rc = torch.randn(E, 2, generator=gen)
    att_rp = torch.randn(E, 3, generator=gen)

    class G:
        def edata(self, _t, name):
            return {"type": etype, "att_rc": att_rc, "att_rp": att_rp}[name]

    out = emb(G())
    d = dict(type=np_(etype), att_rc=np_(att_rc), att_rp=np_(att_rp), out=np_(out))
    for k, p in emb.named_parameters():
        d["p." + k] = np_(p)
    np.savez_compressed(os.path.join(HERE, fname), **d)


# ---------------------------------------------------------------- a4: JSSP (DGL flavour) batching
def _install_dgl_batching_shim():
    """dgl.add_self_loop / dgl.batch / DGLGraph.add_edges for
    jssp/models/agent_observation.py (published DGL semantics: add_self_loop
    appends one (i,i) edge per node after the existing edges with the named edge
    features filled with ``fill_data``; batch concatenates graphs in order,
    offsetting node ids by the cumulative node count)."""
    import dgl

    def add_edges(self, u, v, data=None):
        u = torch.as_tensor(u).long()
        v = torch.as_tensor(v).long()
        n_old = self._src.numel()
        self._src = torch.cat([self._src, u])
        self._dst = torch.cat([self._dst, v])
        for k in list(self.edata.keys()):
            new = data[k] if (data is not None and k in data) else torch.zeros(
                (u.numel(),) + tuple(self.edata[k].shape[1:]), dtype=self.edata[k].dtype)
            self.edata[k] = torch.cat([self.edata[k], torch.as_tensor(new).to(self.edata[k].dtype)])
        assert all(v_.shape[0] == n_old + u.numel() for v_ in self.edata.values())

    def to(self, device):
        return self

    def add_self_loop(g, edge_feat_names=None, fill_data=1.0):
        new = dgl.graph((g._src.clone(), g._dst.clone()), num_nodes=g._n)
        new.ndata = dict(g.ndata)
        new.edata = dict(g.edata)
        loop = torch.arange(g._n)
        data = {k: torch.full((g._n,), fill_data, dtype=new.edata[k].dtype)
                for k in (edge_feat_names or [])}
        new.add_edges(loop, loop, data=data)
        return new

    def batch(graphs):
        off = 0
        srcs, dsts = [], []
        for g in graphs:
            srcs.append(g._src + off)
            dsts.append(g._dst + off)
            off += g._n
        out = dgl.graph((torch.cat(srcs), torch.cat(dsts)), num_nodes=off)
        for k in graphs[0].ndata:
            out.ndata[k] = torch.cat([g.ndata[k] for g in graphs])
        for k in graphs[0].edata:
            out.edata[k] = torch.cat([g.edata[k] for g in graphs])
        out.batch_size = len(graphs)
        out.set_batch_num_nodes = lambda t: setattr(out, "_bnn", t)
        out.set_batch_num_edges = lambda t: setattr(out, "_bne", t)
        out.batch_num_nodes = lambda: out._bnn
        out.batch_num_edges = lambda: out._bne
        return out

    dgl.DGLGraph.add_edges = add_edges
    dgl.DGLGraph.to = to
    dgl.add_self_loop = add_self_loop
    dgl.batch = batch
    # agent_observation.py imports generic.tokengt.utils (laplacian PE helper, unused here)
    stub = types.ModuleType("generic.tokengt.utils")
    stub.get_laplacian_pe_simple = None
    sys.modules.setdefault("generic.tokengt", types.ModuleType("generic.tokengt"))
    sys.modules["generic.tokengt.utils"] = stub


def synthetic_jssp_obs(n_j, n_m, batch, seed, scheduled_frac=0.0, feat_dim=10, pad_nodes=0, pad_edges=0):
    """A padded gym observation dict in the layout of jssp/env/observation.py:83-128:
    precedence edges ordered as jssp/env/state.py:220-227, plus machine-chain edges for
    a seeded random dispatch order (state.py:655-674 appends them one by one)."""
    rng = np.random.default_rng(seed)
    n = n_j * n_m
    feats, eis, nes = [], [], []
    for b in range(batch):
        machines = np.stack([rng.permutation(n_m) for _ in range(n_j)])   # [n_j, n_m]
        f = rng.random((n, feat_dim)).astype(np.float32)
        f[:, 6] = machines.reshape(-1)
        f[:, 1] = np.where(rng.random(n) < 0.3, -1.0, f[:, 1])
        edges = [(j * n_m + i, j * n_m + i + 1) for i in range(n_m - 1) for j in range(n_j)]
        # machine chains: dispatch a random prefix of operations job by job
        n_sched = int(scheduled_frac * n)
        next_op = np.zeros(n_j, dtype=np.int64)
        last_on_machine = {}
        for _ in range(n_sched):
            cand = np.nonzero(next_op < n_m)[0]
            j = int(rng.choice(cand))
            node = j * n_m + int(next_op[j])
            m = int(machines[j, next_op[j]])
            if m in last_on_machine:
                edges.append((last_on_machine[m], node))
            last_on_machine[m] = node
            next_op[j] += 1
        ei = np.array(edges, dtype=np.int64).T
        feats.append(f)
        eis.append(ei)
        nes.append(ei.shape[1])
    max_e = max(nes) + pad_edges
    F = np.zeros((batch, n + pad_nodes, feat_dim), dtype=np.float32)
    EI = np.zeros((batch, 2, max_e), dtype=np.int64)
    for b in range(batch):
        F[b, :n] = feats[b]
        EI[b, :, : nes[b]] = eis[b]
    return {"features": torch.tensor(F), "edge_index": torch.tensor(EI),
            "n_nodes": torch.tensor([n] * batch), "n_edges": torch.tensor(nes)}


def golden_jssp_batch(fname: str, seed: int):
    _install_dgl_batching_shim()
    from jssp.models.agent_observation import AgentObservation
    from jssp.utils.utils import compute_conflicts_cliques

    d = {}
    mid = torch.tensor([0, 1, -1, 1, 0, 2, 2, 0, -1, 1])
    ce, cemid = compute_conflicts_cliques(mid)
    d.update(clq_mid=np_(mid), clq_edges=np_(ce), clq_edges_mid=np_(cemid))
    for tag, conflicts, mie in [("att", "att", False), ("clique", "clique", False),
                                ("cliquemid", "clique", True)]:
        obs = synthetic_jssp_obs(4, 3, 3, seed, scheduled_frac=0.5, pad_nodes=2, pad_edges=3)
        ao = AgentObservation.from_gym_observation(obs, conflicts=conflicts, max_n_machines=3,
                                                   mid_in_edges=mie)
        g = ao.to_graph()
        for k, v in obs.items():
            d[f"{tag}.obs.{k}"] = np_(v)
        d[f"{tag}.src"], d[f"{tag}.dst"] = np_(g._src), np_(g._dst)
        d[f"{tag}.type"] = np_(g.edata["type"])
        d[f"{tag}.feat"] = np_(g.ndata["feat"])
        d[f"{tag}.bnn"], d[f"{tag}.bne"] = np_(g.batch_num_nodes()), np_(g.batch_num_edges())
    np.savez_compressed(os.path.join(HERE, fname), **d)


# ---------------------------------------------------------------- trunk: jssp/models/gnn_dgl.py GnnDGL.forward
def _install_dgl_trunk_shim():
    """What ``jssp/models/gnn_dgl.py`` needs on top of the batching shim: ``dgl.nn.EGATConv`` (:= the in-repo
    restatement ``generic/eegatconv.py`` EEGATConv with ``edge_scoring=False``, SURVEY F3: same math as DGL's
    layer), ``DGLGraph.add_nodes`` (published DGL semantics: new nodes get ids after the existing ones, the
    given node features are appended, missing ones zero-filled), and placeholders for the layer classes the
    file imports but the gat/gatv2 branch never instantiates."""
    import dgl
    from generic.eegatconv import EEGATConv

    class EGATConv(EEGATConv):
        def __init__(self, in_node_feats, in_edge_feats, out_node_feats, out_edge_feats, num_heads, bias=True):
            super().__init__(in_node_feats, in_edge_feats, out_node_feats, out_edge_feats, num_heads, bias=bias,
                             edge_scoring=False)

        def forward(self, graph, nfeats, efeats, edge_weight=None, get_attention=False):
            return super().forward(graph, nfeats, efeats, None, edge_weight, get_attention)

    def add_nodes(self, num, data=None):
        self._n += num
        for k in list(self.ndata.keys()):
            new = data[k] if (data is not None and k in data) else torch.zeros(
                (num,) + tuple(self.ndata[k].shape[1:]), dtype=self.ndata[k].dtype)
            self.ndata[k] = torch.cat([self.ndata[k], torch.as_tensor(new).to(self.ndata[k].dtype)])

    dgl.DGLGraph.add_nodes = add_nodes
    dgl.nn.EGATConv = EGATConv
    for name in ("DGNConv", "GCN2Conv", "GINEConv", "PNAConv"):
        setattr(dgl.nn, name, type(name, (), {}))
    dgl.LaplacianPE = type("LaplacianPE", (), {})


def golden_gnn_dgl(fname: str, seed: int):
    """The reference's own ``GnnDGL.forward`` (jssp/models/gnn_dgl.py:245-431) executed on the padded gym
    observation: batching, pool / machine node wiring, embedder, L x (EGATConv -> MLP, residual), concat, pool
    slice, pad, final cat.  float32 (the reference module's dtype)."""
    _install_dgl_batching_shim()
    _install_dgl_trunk_shim()
    from jssp.models.gnn_dgl import GnnDGL

    d = {}
    n_j, n_m, B, Fdim, Hd, H, L = 4, 3, 3, 10, 8, 2, 2
    for tag, conflicts, pooling, mie, pad_nodes in [("clique", "clique", "learn", False, 0),
                                                   ("cliquepad", "clique", "learn", True, 3),
                                                   ("node", "node", "learn", True, 0),
                                                   ("att_max", "att", "max", False, 2)]:
        obs = synthetic_jssp_obs(n_j, n_m, B, seed, scheduled_frac=0.5, feat_dim=Fdim, pad_nodes=2, pad_edges=3)
        torch.manual_seed(seed)
        gnn = GnnDGL(Fdim, "gatv2", pooling, False, n_j * n_m + pad_nodes, n_m, 3, L, Hd, "gelu", H,
                     residual=True, normalize=False, conflicts=conflicts, no_tct=False, mid_in_edges=mie)
        with torch.no_grad():
            for fe in gnn.features_extractors:
                fe.bias.copy_(torch.randn(fe.bias.shape) * 0.1)
        ret = gnn(obs)
        gen = torch.Generator().manual_seed(seed + 1)
        R = torch.randn(ret.shape, generator=gen)
        names = [k for k, _ in gnn.named_parameters()]
        params = [p for _, p in gnn.named_parameters()]
        gr = grads_of((ret * R).sum(), params)
        for k, v in obs.items():
            d[f"{tag}.obs.{k}"] = np_(v)
        d[f"{tag}.ret"], d[f"{tag}.R"] = np_(ret), np_(R)
        d[f"{tag}.cfg"] = np.array([n_j, n_m, B, Fdim, Hd, H, L, n_j * n_m + pad_nodes, int(mie)])
        d[f"{tag}.conflicts"], d[f"{tag}.pooling"] = np.array(conflicts), np.array(pooling)
        for k, p, g_ in zip(names, params, gr):
            d[f"{tag}.p.{k}"] = np_(p)
            d[f"{tag}.g.{k}"] = np_(g_)
    np.savez_compressed(os.path.join(HERE, fname), **d)


# ---------------------------------------------------------------- a2: GraphConv -> PyG GATv2Conv
def golden_gatv2(fname: str, seed: int):
    """``psp/graph/graph_conv.py`` GraphConv(pyg=True) executed verbatim: its constructor call (:33-40), ``forward``
    (:51-53) and ``forward_nog`` (:59-60) - on ``oracle/shims/torch_geometric`` GATv2Conv, a RESTATEMENT of the
    third-party layer (parity unpinned for the layer itself, see the class docstring)."""
    from psp.graph.graph_conv import GraphConv

    gen = torch.Generator().manual_seed(seed)
    N, E0, Fin, C, H = 12, 36, 6, 5, 3
    src, dst = rand_graph(gen, N, E0, self_loops=False)
    dst[dst == 7] = 3                                   # node 7 has no in-edge: output = bias
    ei = torch.stack([src, dst])
    torch.manual_seed(seed)
    layer = GraphConv(Fin, C, H, pyg=True).double()
    assert layer.gatv3 is False
    with torch.no_grad():
        for p_ in (layer.conv.lin_l.bias, layer.conv.lin_r.bias, layer.conv.bias):
            p_.copy_(torch.randn(p_.shape, generator=gen, dtype=torch.float64) * 0.1)
    x = torch.randn(N, Fin, generator=gen, dtype=torch.float64, requires_grad=True)
    ea = torch.randn(E0, Fin, generator=gen, dtype=torch.float64, requires_grad=True)
    g = types.SimpleNamespace(edge_index=ei)
    out, none_ = layer(g, x, ea)
    assert none_ is None
    out_nog, _ = layer.forward_nog(x, ei, ea)
    assert torch.equal(out, out_nog)
    out_noedge, _ = layer.forward_nog(x, ei, None)      # psp/models/gnn_hier.py:69,115
    _, (_, alpha) = layer.conv(x, ei, ea, return_attention_weights=True)
    R = torch.randn(out.shape, generator=gen, dtype=torch.float64)
    names = [k for k, _ in layer.named_parameters()]
    params = [p_ for _, p_ in layer.named_parameters()]
    gr = grads_of((out * R).sum(), params + [x, ea])
    d = dict(edge_index=np_(ei), x=np_(x), ea=np_(ea), R=np_(R), out=np_(out), out_noedge=np_(out_noedge),
             alpha=np_(alpha), gx=np_(gr[-2]), gea=np_(gr[-1]), heads=np.int64(H))
    for k, p_, gp in zip(names, params, gr):
        d["p." + k] = np_(p_)
        d["g." + k] = np_(gp)
    np.savez_compressed(os.path.join(HERE, fname), **d)


# ---------------------------------------------------------------- a5: PSP / PyG-flavour batching (Agent.preprocess)
def _install_psp_stubs():
    """Modules psp/models/agent.py imports that need third-party wheels absent here and are NOT on the path
    exercised (hierarchical / TokenGT extractors, the HeteroData graph factory): empty placeholders."""
    for name, attrs in (("psp.models.gnn_hier", ["GnnHier"]), ("generic.tokenGT", ["TokenGT"]),
                        ("psp.graph.graph_factory", ["GraphFactory"])):
        if name not in sys.modules:
            m = types.ModuleType(name)
            for a in attrs:
                setattr(m, a, type(a, (), {}))
            sys.modules[name] = m


def synthetic_psp_env_graph(problem, rng, n_rp=0, with_rp_store=False):
    """The graph an env worker hands over (psp/env/gstate.py:169-183,544-546 on the stand-in container): ``prec``
    edges in ``problem_edges`` order (gstate.py:98-103), the node fields ``rewire_internal`` concatenates
    (agent_graph_observation.py:66-86) with seeded random values, ``resources`` normalised by availability
    (gstate.py:260-277), ``res_cal`` (:332-341), optionally ``n_rp`` resource-priority edges with their 4 attributes
    (gstate.py:891-934)."""
    from psp_graph import StandInGraph
    job_modes, m = [], 0
    for j in problem.n_modes_per_job:
        job_modes.append(list(range(m, m + j)))
        m += j
    n = m
    edges = [(o, d) for k, succ in enumerate(problem.successors_id) for sj in succ
             for o in job_modes[k] for d in job_modes[sj]]
    g = StandInGraph(edges, n)
    f32 = lambda a: torch.tensor(np.asarray(a), dtype=torch.float32)   # noqa: E731
    g.set_ndata("job", torch.tensor([k for k, jm in enumerate(job_modes) for _ in jm], dtype=torch.int32))
    g.set_ndata("durations", f32(rng.random((n, 3))))
    g.set_ndata("affected", f32(rng.random(n) < 0.3))
    g.set_ndata("selectable", f32(rng.random(n) < 0.3))
    typ = np.zeros(n, np.float32)
    typ[0], typ[-1] = -1, 1
    g.set_ndata("type", f32(typ))
    g.set_ndata("normalized_durations", f32(rng.random((n, 3))))
    g.set_ndata("normalized_tct", f32(rng.random((n, 3))))
    g.set_ndata("normalized_tardiness", f32(rng.random((n, 3))))
    g.set_ndata("has_due_date", f32(rng.random(n) < 0.2))
    g.set_ndata("normalized_due_dates", f32(rng.random(n)))
    g.set_ndata("past", torch.tensor(rng.random(n) < 0.2))
    res = np.array([c for per_job in problem.resource_cons for c in per_job], dtype=np.float32)
    res = res / np.array(problem.resource_availabilities, dtype=np.float32)[None, :]
    g.set_ndata("resources", f32(res))
    g.set_global_data("res_cal", torch.tensor([(0, 0, 0)] * problem.n_resources, dtype=torch.float32))
    if n_rp:
        s = torch.tensor(rng.integers(0, n, n_rp))
        d = torch.tensor(rng.integers(0, n, n_rp))
        r = f32(np.concatenate([rng.integers(0, problem.n_resources, (n_rp, 1)), rng.random((n_rp, 3))], axis=1))
        g.add_edges(s, d, etype="rp", data={"r": r})
    elif with_rp_store:
        g.num_edges("rp")
    return g


def golden_psp_batch(fname: str, seed: int):
    """``Agent._obs_as_tensor_graph`` (rewire_internal) and ``Agent.preprocess`` of psp/models/agent.py EXECUTED
    VERBATIM (unbound, on a namespace carrying the specification fields they read) together with everything they
    call - AgentGraphObservation, rewire, node_conflicts, learned_graph_pool, vnode, homogeneous_edges,
    compute_resources_graph_torch - on the stand-in graph container (oracle/shims/psp_graph.py)."""
    _install_psp_stubs()
    from psp.models.agent import Agent as PspAgent
    from psp.utils.loaders import PSPLoader
    from psp.utils.taillard_rcpsp import TaillardRcpsp
    import contextlib
    import io

    with contextlib.redirect_stdout(io.StringIO()):
        p272 = PSPLoader().load_single("/root/reference/instances/psp/272/272.sm")
    tai43 = TaillardRcpsp(0, 4, 3, seed=3)
    tai33 = TaillardRcpsp(1, 3, 3, seed=4)
    d = {}
    cases = [
        # tag, conflicts, pooling, vnode, self loops, [(problem, n_rp, with_rp_store)]
        ("node272", "node", "learn", False, True, [(p272, 0, False), (p272, 5, False), (p272, 2, False)]),
        ("clique272", "clique", "learn", False, True, [(p272, 3, False), (p272, 0, False)]),
        # vnode=True is not runnable in the reference: rewiring.py:288-290 hands Python lists to add_edges, whose
        # torch.stack (pyg_graph.py:168) raises TypeError
        ("nodeinv", "node", "learninv", False, True, [(tai43, 0, True), (tai43, 4, False)]),
        ("ragged", "node", "learn", False, True, [(p272, 2, False), (tai33, 0, False), (tai43, 3, False)]),
        ("clique_noself", "clique", "max", False, False, [(tai43, 0, False), (tai33, 2, False)]),
        ("single", "node", "learn", False, True, [(tai33, 1, False)]),
    ]
    for tag, conflicts, pooling, vnode, self_loops, graphs in cases:
        rng = np.random.default_rng(seed)
        env_graphs = [synthetic_psp_env_graph(pb, rng, n_rp, st) for pb, n_rp, st in graphs]
        ns = types.SimpleNamespace(
            agent_specification=types.SimpleNamespace(conflicts=conflicts, graph_pooling=pooling, vnode=vnode,
                                                      rwpe_k=0),
            env_specification=types.SimpleNamespace(factored_rp=False, add_rp_edges="frontier_strict",
                                                    max_n_resources=max(pb.n_resources for pb, _, _ in graphs),
                                                    remove_past_prec=self_loops, observe_subgraph=False,
                                                    n_features=16),
            graphobs=True, rwpe_cache=None)
        # wire-level inputs: what the env worker's graph holds (before rewire_internal)
        d[f"{tag}.n_graphs"] = np.int64(len(env_graphs))
        d[f"{tag}.opts"] = np.array([conflicts, pooling, str(int(vnode)), str(int(self_loops))])
        for i, g in enumerate(env_graphs):
            d[f"{tag}.g{i}.prec"] = np_(g._graph.edges["prec"]["edge_index"])
            d[f"{tag}.g{i}.resources"] = np_(g.ndata("resources"))
            d[f"{tag}.g{i}.res_cal"] = np_(g.global_data("res_cal"))
            rp = g._graph.edges.get("rp")
            has = rp is not None and "edge_index" in rp
            d[f"{tag}.g{i}.rp"] = np_(rp["edge_index"]) if has else np.zeros((2, 0), np.int64)
            d[f"{tag}.g{i}.rp_r"] = np_(rp["r"]) if has else np.zeros((0, 4), np.float32)
            d[f"{tag}.g{i}.stores_env"] = np.array(list(g._graph.edges.keys()))
        obs = PspAgent._obs_as_tensor_graph(ns, [g.clone() for g in env_graphs])      # alg/ppo.py:144,236
        for i, g in enumerate(obs):
            d[f"{tag}.g{i}.feat"] = np_(g.ndata("feat"))                               # [n, 16]
            d[f"{tag}.g{i}.stores"] = np.array(list(g._graph.edges.keys()))
        (g, batch_size, num_nodes, n_nodes, poolnodes, resource_nodes, vnodes, res_cal_id, pe) = \
            PspAgent.preprocess(ns, obs)
        assert pe is None
        h = g._graph
        d[f"{tag}.out.edge_index"] = np_(h["edge_index"])
        for k in ("type", "rid", "att_rc", "att_rp", "feat"):
            d[f"{tag}.out.{k}"] = np_(h[k])
        d[f"{tag}.out.store_order"] = np.array(h.store_order)
        d[f"{tag}.out.batch_size"], d[f"{tag}.out.num_nodes"] = np.int64(batch_size), np.int64(num_nodes)
        d[f"{tag}.out.n_nodes"] = np.atleast_1d(np.asarray(n_nodes, dtype=np.int64))
        d[f"{tag}.out.poolnodes"] = np.asarray([] if poolnodes is None else poolnodes, dtype=np.int64)
        d[f"{tag}.out.resource_nodes"] = (np.zeros(0, np.int64) if resource_nodes is None
                                          else np_(resource_nodes).astype(np.int64))
        d[f"{tag}.out.vnodes"] = np.asarray([] if vnodes is None else vnodes, dtype=np.int64)
        d[f"{tag}.out.res_cal_id"] = np_(res_cal_id)
    np.savez_compressed(os.path.join(HERE, fname), **d)


# ---------------------------------------------------------------- N3: GnnMP / GnnFlat trunk (PyG flavour)
def golden_gnn_mp(fname: str, seed: int):
    """psp/models/gnn_mp.py GnnMP.forward (:318-507) and psp/models/gnn_flat.py GnnFlat.forward (:151-259) EXECUTED
    VERBATIM on the batches ``Agent.preprocess`` builds (stand-in container) with the layer stack on the PyG shim
    (GraphConv -> GATv2Conv restatement, PspEdgeEmbedder and MLP are the reference's own).  ``torch.compile`` is
    kept (so the ``_orig_mod.`` infix of the state-dict keys is the reference's) but pointed at the ``eager``
    backend: no Inductor / Triton in this container, same arithmetic."""
    _install_psp_stubs()
    from psp.models.agent import Agent as PspAgent
    from psp.models.gnn_mp import GnnMP
    from psp.utils.loaders import PSPLoader
    from psp.utils.taillard_rcpsp import TaillardRcpsp
    import contextlib
    import io

    with contextlib.redirect_stdout(io.StringIO()):
        p272 = PSPLoader().load_single("/root/reference/instances/psp/272/272.sm")
    tai33 = TaillardRcpsp(1, 3, 3, seed=4)
    orig_compile = torch.compile
    torch.compile = lambda m, **kw: orig_compile(m, backend="eager", dynamic=kw.get("dynamic"))
    d = {}
    try:
        for tag, conflicts, pooling, graphs, max_n_nodes in [
                ("node272", "node", "learn", [(p272, 0), (p272, 5), (p272, 2)], 20),
                ("clique272", "clique", "learn", [(p272, 3), (p272, 0)], 18),
                ("ragged_max", "node", "max", [(p272, 2), (tai33, 0)], 19),
                ("single", "node", "learn", [(tai33, 1)], 13)]:
            rng = np.random.default_rng(seed)
            env_graphs = [synthetic_psp_env_graph(pb, rng, n_rp, False) for pb, n_rp in graphs]
            max_res = max(pb.n_resources for pb, _ in graphs)
            ns = types.SimpleNamespace(
                agent_specification=types.SimpleNamespace(conflicts=conflicts, graph_pooling=pooling, vnode=False,
                                                          rwpe_k=0),
                env_specification=types.SimpleNamespace(factored_rp=False, add_rp_edges="frontier_strict",
                                                        max_n_resources=max_res, remove_past_prec=True,
                                                        observe_subgraph=False, n_features=16),
                graphobs=True, rwpe_cache=None)
            d[f"{tag}.n_graphs"] = np.int64(len(env_graphs))
            d[f"{tag}.opts"] = np.array([conflicts, pooling, "0", "1"])
            obs = PspAgent._obs_as_tensor_graph(ns, [g.clone() for g in env_graphs])
            for i, (g, ge) in enumerate(zip(obs, env_graphs)):
                d[f"{tag}.g{i}.feat"] = np_(g.ndata("feat"))
                d[f"{tag}.g{i}.prec"] = np_(ge._graph.edges["prec"]["edge_index"])
                d[f"{tag}.g{i}.resources"] = np_(ge.ndata("resources"))
                d[f"{tag}.g{i}.res_cal"] = np_(ge.global_data("res_cal")) + 1.0 + i   # non-trivial calendar ids
                ge.set_global_data("res_cal", torch.tensor(d[f"{tag}.g{i}.res_cal"]))
                g.set_global_data("res_cal", torch.tensor(d[f"{tag}.g{i}.res_cal"]))
                rp = ge._graph.edges.get("rp")
                has = rp is not None and "edge_index" in rp
                d[f"{tag}.g{i}.rp"] = np_(rp["edge_index"]) if has else np.zeros((2, 0), np.int64)
                d[f"{tag}.g{i}.rp_r"] = np_(rp["r"]) if has else np.zeros((0, 4), np.float32)
                d[f"{tag}.g{i}.stores_env"] = np.array(list(ge._graph.edges.keys()))
            pre = PspAgent.preprocess(ns, obs)
            Hd, H, L = 8, 2, 2
            torch.manual_seed(seed)
            gnn = GnnMP(input_dim_features_extractor=16, graph_pooling=pooling, max_n_nodes=max_n_nodes,
                        max_n_resources=max_res, n_mlp_layers_features_extractor=3, n_layers_features_extractor=L,
                        hidden_dim_features_extractor=Hd, activation_features_extractor="gelu",
                        n_attention_heads=H, residual=True, normalize=False, conflicts=conflicts,
                        edge_embedding_flavor="sum", layer_pooling="all", factored_rp=False,
                        add_rp_edges="frontier_strict", add_self_loops=True, vnode=False,
                        update_edge_features=False, update_edge_features_pe=False, rwpe_k=0, rwpe_h=0,
                        rwpe_cache=None, graphobs=True, shared_conv=False, pyg=True, hierarchical=False,
                        tokengt=False, checkpoint=1, dual_net=False)
            with torch.no_grad():
                for fe in gnn.gnn._orig_mod.features_extractors:
                    for p_ in (fe.conv.lin_l.bias, fe.conv.lin_r.bias, fe.conv.bias):
                        p_.copy_(torch.randn(p_.shape) * 0.1)
            node_features, graph_embedding = gnn(pre)
            gen = torch.Generator().manual_seed(seed + 1)
            R1 = torch.randn(node_features.shape, generator=gen)
            R2 = torch.randn(graph_embedding.shape, generator=gen)
            names = [k for k, _ in gnn.named_parameters()]
            params = [p_ for _, p_ in gnn.named_parameters()]
            gr = grads_of((node_features * R1).sum() + (graph_embedding * R2).sum(), params)
            d[f"{tag}.node_features"], d[f"{tag}.graph_embedding"] = np_(node_features), np_(graph_embedding)
            d[f"{tag}.R1"], d[f"{tag}.R2"] = np_(R1), np_(R2)
            d[f"{tag}.cfg"] = np.array([Hd, H, L, max_n_nodes, max_res])
            d[f"{tag}.state_keys"] = np.array(list(gnn.state_dict().keys()))
            for k, p_, g_ in zip(names, params, gr):
                d[f"{tag}.p.{k}"] = np_(p_)
                d[f"{tag}.g.{k}"] = np_(g_)
    finally:
        torch.compile = orig_compile
    np.savez_compressed(os.path.join(HERE, fname), **d)


# ---------------------------------------------------------------- N4: GAE + PPO minibatch loss
def _reference_lines(path, first, last):
    """Source lines first..last (1-based, inclusive) of a reference file, dedented: code that sits in the middle
    of a long method (PPO.collect_rollouts / PPO.train, module not importable without gymnasium) is executed from
    the reference tree itself, in a namespace that supplies the variables the method has in scope."""
    import textwrap
    with open(os.path.join("/root/reference", path)) as f:
        lines = f.readlines()[first - 1:last]
    return textwrap.dedent("".join(lines))


def golden_ppo(fname: str, seed: int):
    gen = torch.Generator().manual_seed(seed)
    out = {}
    # ---- GAE: alg/ppo.py:276-299
    src = _reference_lines("alg/ppo.py", 276, 299)
    assert src.startswith("with torch.no_grad():") and "lastgaelam" in src and "returns = advantages + values" in src
    code = compile(src, "/root/reference/alg/ppo.py:276-299", "exec")
    for tag, (T, E, R, gamma, lam, sigma) in dict(a=(7, 3, 1, 1.0, 1.0, 1.0), b=(40, 5, 2, 0.99, 0.95, 1.7),
                                                  c=(1, 4, 1, 0.9, 0.8, 1.0)).items():
        rewards, values = torch.randn(T, E, R, generator=gen), torch.randn(T, E, R, generator=gen)
        dones = (torch.rand(T, E, generator=gen) < 0.2).float()
        next_value, next_done = torch.randn(E, R, generator=gen), (torch.rand(E, generator=gen) < 0.3).float()
        ns = dict(torch=torch, rewards=rewards, values=values, dones=dones, next_value=next_value,
                  next_done=next_done, sigma=sigma,
                  self=types.SimpleNamespace(num_steps=T, gamma=gamma, gae_lambda=lam))
        exec(code, ns)
        out.update({f"gae.{tag}.rewards": np_(rewards), f"gae.{tag}.values": np_(values), f"gae.{tag}.dones": np_(dones),
                    f"gae.{tag}.next_value": np_(next_value), f"gae.{tag}.next_done": np_(next_done),
                    f"gae.{tag}.cfg": np.array([gamma, lam, sigma]), f"gae.{tag}.advantages": np_(ns["advantages"]),
                    f"gae.{tag}.returns": np_(ns["returns"])})
    # ---- minibatch loss: alg/ppo.py:582-705
    src = _reference_lines("alg/ppo.py", 582, 705)
    assert src.startswith("logratio = newlogprob") and src.rstrip().endswith(")") and "uniform_loss" in src
    code = compile(src, "/root/reference/alg/ppo.py:582-705", "exec")
    cases = dict(
        default=dict(B=64, clip_coef=0.2, norm_adv=True, clip_vloss=False, critic_loss="l2", ent_coef=0.0, vf_coef=0.5, update=3),
        entropy_l1=dict(B=37, clip_coef=0.1, norm_adv=True, clip_vloss=False, critic_loss="l1", ent_coef=0.05, vf_coef=2.0, update=2),
        no_norm=dict(B=16, clip_coef=0.3, norm_adv=False, clip_vloss=False, critic_loss="l2", ent_coef=0.01, vf_coef=1.0, update=1),
        espo=dict(B=33, clip_coef=None, norm_adv=True, clip_vloss=False, critic_loss="l2", ent_coef=0.02, vf_coef=0.5, update=5),
        vclip_l2=dict(B=50, clip_coef=0.2, norm_adv=True, clip_vloss=True, critic_loss="l2", ent_coef=0.0, vf_coef=0.5, update=2),
        vclip_sl1=dict(B=41, clip_coef=0.15, norm_adv=True, clip_vloss=True, critic_loss="sl1", ent_coef=0.03, vf_coef=0.7, update=2),
        warmup=dict(B=24, clip_coef=0.2, norm_adv=True, clip_vloss=False, critic_loss="l2", ent_coef=0.0, vf_coef=0.5, update=0),
        single=dict(B=1, clip_coef=0.2, norm_adv=True, clip_vloss=False, critic_loss="l2", ent_coef=0.1, vf_coef=0.5, update=4),
    )
    for tag, c in cases.items():
        B, n_act = c["B"], 9
        newlogprob = (-torch.rand(B, generator=gen) * 3).requires_grad_(True)
        old = (newlogprob.detach() + 0.4 * torch.randn(B, generator=gen))
        entropy = torch.rand(B, generator=gen).requires_grad_(True)
        newvalue = torch.randn(B, 1, generator=gen).requires_grad_(True)
        returns, old_values = torch.randn(B, 1, generator=gen), torch.randn(B, 1, generator=gen)
        adv = torch.randn(B, 1, generator=gen)
        logits = torch.randn(B, n_act, generator=gen).requires_grad_(True)
        unmasked = torch.distributions.Categorical(logits=logits)
        spec = types.SimpleNamespace(two_hot=None, hl_gauss=None, symlog=False)
        agent = types.SimpleNamespace(agent_specification=spec, aggregate_reward=lambda r: r.squeeze(1) * 1.0)
        self_ = types.SimpleNamespace(clip_coef=c["clip_coef"], target_kl=0.04, norm_adv=c["norm_adv"],
                                      critic_loss=c["critic_loss"], clip_vloss=c["clip_vloss"],
                                      ent_coef=c["ent_coef"], vf_coef=c["vf_coef"])
        ns = dict(torch=torch, nn=torch.nn, np=np, self=self_, agent=agent, newlogprob=newlogprob,
                  batched_logprobs=old, train_device="cpu", approx_kl_divs=[], clipfracs=[],
                  approx_kl_divs_on_epoch=[], batched_advantages=adv, max_espo_dev=0, newvalue=newvalue,
                  batched_returns=returns, batched_values=old_values, entropy=entropy, update=c["update"],
                  unmasked_distrib=unmasked)
        if c["critic_loss"] == "sl1":
            # :632-641 define the unclipped term for l2 / l1 only; give it the value :679-684 use for the clipped one
            ns["v_loss_unclipped"] = torch.nn.functional.smooth_l1_loss(newvalue, returns)
        exec(code, ns)
        loss = ns["loss"]
        g = torch.autograd.grad(loss, [newlogprob, entropy, newvalue, logits], allow_unused=True)
        z = lambda t, like: np_(torch.zeros_like(like) if t is None else t)
        out.update({f"loss.{tag}.logprob": np_(newlogprob), f"loss.{tag}.old_logprob": np_(old),
                    f"loss.{tag}.entropy": np_(entropy), f"loss.{tag}.value": np_(newvalue),
                    f"loss.{tag}.returns": np_(returns), f"loss.{tag}.old_values": np_(old_values),
                    f"loss.{tag}.advantages": np_(adv), f"loss.{tag}.logits": np_(logits),
                    f"loss.{tag}.cfg": np.array([str(c[k]) for k in ("clip_coef", "norm_adv", "clip_vloss", "critic_loss",
                                                                     "ent_coef", "vf_coef", "update")]),
                    f"loss.{tag}.loss": np_(loss), f"loss.{tag}.pg_loss": np_(ns["pg_loss"]),
                    f"loss.{tag}.v_loss": np_(ns["v_loss"]), f"loss.{tag}.entropy_loss": np_(ns["entropy_loss"]),
                    f"loss.{tag}.approx_kl": np.array(ns["approx_kl_divs"][-1]),
                    f"loss.{tag}.clipfrac": np.array(ns["clipfracs"][-1] if ns["clipfracs"] else 0.0),
                    f"loss.{tag}.g_logprob": z(g[0], newlogprob), f"loss.{tag}.g_entropy": z(g[1], entropy),
                    f"loss.{tag}.g_value": z(g[2], newvalue), f"loss.{tag}.g_logits": z(g[3], logits)})
    out["loss.tags"] = np.array(list(cases))
    np.savez_compressed(os.path.join(HERE, fname), **out)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "ppo":
        golden_ppo("ppo.npz", 22)
        sys.exit(0)
    golden_eegat(False, "eegat.npz", 11)
    golden_eegat(True, "eegat_edge_scoring.npz", 12)
    golden_gatv3("gatv3.npz", 13)
    golden_head("head.npz", 14)
    golden_mlp("mlp.npz", 15)
    golden_edge_embedder("edge_embedder.npz", 16)
    golden_jssp_batch("jssp_batch.npz", 17)
    golden_gnn_dgl("gnn_dgl.npz", 18)
    golden_gatv2("gatv2.npz", 19)
    golden_psp_batch("psp_batch.npz", 20)
    golden_gnn_mp("gnn_mp.npz", 21)
    golden_ppo("ppo.npz", 22)
    print("golden fixtures written to", HERE)
