"""The oracle (oracle/) against the golden vectors produced by executing the
reference (tests/golden/make_golden.py).  CPU only."""
import os

import numpy as np
import pytest
import torch

from oracle import batch_oracle, gat_oracle, head_oracle

T = lambda a: torch.from_numpy(np.asarray(a))


def load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name), allow_pickle=False)


def test_egat_forward_backward_matches_reference(golden_dir):
    g = load(golden_dir, "eegat.npz")
    x = T(g["x"]).requires_grad_(True)
    ef = T(g["ef"]).requires_grad_(True)
    P = {k[2:]: T(g[k]).requires_grad_(True) for k in g.files if k.startswith("p.")}
    h, f, a = gat_oracle.egat_forward(x, T(g["src"]), T(g["dst"]), ef, P["fc_ni.weight"],
                                      P["fc_nj.weight"], P["fc_fij.weight"],
                                      P["fc_node_src.weight"], P["attn"], P["bias"])
    np.testing.assert_allclose(h.detach().numpy(), g["h"], rtol=1e-12, atol=1e-13)
    np.testing.assert_allclose(f.detach().numpy(), g["f"], rtol=1e-12, atol=1e-13)
    np.testing.assert_allclose(a.detach().numpy(), g["a"], rtol=1e-12, atol=1e-13)
    names = list(P)
    grads = torch.autograd.grad((h * T(g["R"])).sum(), [P[k] for k in names] + [x, ef])
    for k, gr in zip(names, grads):
        np.testing.assert_allclose(gr.numpy(), g["g." + k], rtol=1e-10, atol=1e-12, err_msg=k)
    np.testing.assert_allclose(grads[-2].numpy(), g["gx"], rtol=1e-10, atol=1e-12)
    np.testing.assert_allclose(grads[-1].numpy(), g["gef"], rtol=1e-10, atol=1e-12)


def test_egat_zero_in_degree_raises():
    x = torch.randn(3, 4, dtype=torch.float64)
    W = torch.randn(6, 4, dtype=torch.float64)
    with pytest.raises(ValueError):
        gat_oracle.egat_forward(x, torch.tensor([0, 1]), torch.tensor([1, 1]),
                                torch.randn(2, 4, dtype=torch.float64), W, W, W, W,
                                torch.randn(1, 2, 3, dtype=torch.float64), None)


def test_eegat_edge_scoring_matches_reference(golden_dir):
    g = load(golden_dir, "eegat_edge_scoring.npz")
    P = {k[2:]: T(g[k]) for k in g.files if k.startswith("p.")}
    h, f, e_out, a = gat_oracle.eegat_forward_edge_scoring(
        T(g["x"]), T(g["src"]), T(g["dst"]), T(g["ef"]), T(g["efe"]), P)
    for got, key in ((h, "h"), (f, "f"), (e_out, "e_out"), (a, "a")):
        np.testing.assert_allclose(got.numpy(), g[key], rtol=1e-12, atol=1e-13, err_msg=key)


def test_gatv3_matches_reference(golden_dir):
    g = load(golden_dir, "gatv3.npz")
    P = {k[2:]: T(g[k]).requires_grad_(True) for k in g.files if k.startswith("p.")}
    assert "bias" not in P                      # reference quirk: no output bias
    x = T(g["x"]).requires_grad_(True)
    out, alpha, ei = gat_oracle.gatv3_forward(
        x, T(g["edge_index"]), P["lin_l.weight"], P["lin_l.bias"], P["lin_r.weight"],
        P["lin_r.bias"], P["context_attention_layer.weight"], P["context_attention_layer.bias"],
        None, int(g["heads"]))
    assert np.array_equal(ei.numpy(), g["edge_index_out"])
    np.testing.assert_allclose(out.detach().numpy(), g["out"], rtol=1e-12, atol=1e-13)
    np.testing.assert_allclose(alpha.detach().numpy(), g["alpha"], rtol=1e-12, atol=1e-13)
    names = [k for k in P if k != "lin_edge.weight"]
    grads = torch.autograd.grad((out * T(g["R"])).sum(), [P[k] for k in names] + [x])
    for k, gr in zip(names, grads):
        np.testing.assert_allclose(gr.numpy(), g["g." + k], rtol=1e-10, atol=1e-12, err_msg=k)
    np.testing.assert_allclose(grads[-1].numpy(), g["gx"], rtol=1e-10, atol=1e-12)
    assert not g["g.lin_edge.weight"].any()     # lin_edge is unused by message()


def test_gatv2_oracle_self_consistency():
    """GATv2 is 'parity unpinned' (third-party layer).  Check the restatement against a
    per-node python loop of the GATv2 definition and with gradcheck."""
    torch.manual_seed(0)
    N, E, Fin, C, H = 7, 25, 5, 4, 3
    layer = gat_oracle.OracleGATv2Conv(Fin, C, H, edge_dim=Fin).double()
    with torch.no_grad():
        for p in layer.parameters():
            p.add_(0.1 * torch.randn_like(p))
    x = torch.randn(N, Fin, dtype=torch.float64)
    ei = torch.randint(0, N, (2, E))
    ea = torch.randn(E, Fin, dtype=torch.float64)
    out = layer(x, ei, ea)
    xl = (x @ layer.lin_l.weight.t() + layer.lin_l.bias).view(N, H, C)
    xr = (x @ layer.lin_r.weight.t() + layer.lin_r.bias).view(N, H, C)
    pe = (ea @ layer.lin_edge.weight.t()).view(E, H, C)
    ref = torch.zeros(N, H, C, dtype=torch.float64)
    for v in range(N):
        es = [e for e in range(E) if int(ei[1, e]) == v]
        if not es:
            continue
        for h in range(H):
            s = torch.stack([(torch.nn.functional.leaky_relu(
                xl[int(ei[0, e]), h] + xr[v, h] + pe[e, h], 0.2) * layer.att[0, h]).sum() for e in es])
            w = torch.exp(s - s.max())
            w = w / (w.sum() + 1e-16)
            for k, e in enumerate(es):
                ref[v, h] += w[k] * xl[int(ei[0, e]), h]
    ref = ref.reshape(N, H * C) + layer.bias
    np.testing.assert_allclose(out.detach().numpy(), ref.detach().numpy(), rtol=1e-12, atol=1e-13)
    xg = x.clone().requires_grad_(True)
    assert torch.autograd.gradcheck(lambda t: layer(t, ei, ea), (xg,), eps=1e-6, atol=1e-6)
    # isolated destination -> bias only (SURVEY 8b 'Errors')
    out2 = layer(x, torch.tensor([[0, 1], [1, 1]]), ea[:2])
    np.testing.assert_allclose(out2[2].detach().numpy(), layer.bias.detach().numpy())


def test_head_matches_reference(golden_dir):
    g = load(golden_dir, "head.npz")
    nf = T(g["node_features"]).requires_grad_(True)
    logits = (nf @ T(g["w_action"]).t() + T(g["b_action"])).squeeze(-1)
    masks = T(g["masks"])
    a, lp, ent, _ = head_oracle.action_and_logprob(logits, None, masks, deterministic=True)
    assert np.array_equal(a.numpy(), g["det_action"])
    np.testing.assert_array_equal(lp.detach().numpy(), g["det_logprob"])
    np.testing.assert_array_equal(ent.detach().numpy(), g["det_entropy"])
    a2, lp2, ent2, _ = head_oracle.action_and_logprob(logits, T(g["action_in"]), masks)
    np.testing.assert_array_equal(lp2.detach().numpy(), g["eval_logprob"])
    np.testing.assert_array_equal(ent2.detach().numpy(), g["eval_entropy"])
    (gnf,) = torch.autograd.grad((lp2 * T(g["wl"])).sum() + (ent2 * T(g["we"])).sum(), [nf])
    np.testing.assert_allclose(gnf.numpy(), g["g_nf"], rtol=1e-6, atol=1e-7)
    _, lp3, ent3, _ = head_oracle.action_and_logprob(logits, T(g["action_in"]), None)
    np.testing.assert_array_equal(lp3.detach().numpy(), g["nomask_logprob"])
    np.testing.assert_allclose(ent3.detach().numpy(), g["nomask_entropy"], rtol=1e-6)
    torch.manual_seed(1234)
    a4, lp4, _, _ = head_oracle.action_and_logprob(logits.detach(), None, masks, deterministic=False)
    assert np.array_equal(a4.numpy(), g["sample_action"])
    np.testing.assert_array_equal(lp4.numpy(), g["sample_logprob"])
    np.testing.assert_array_equal(head_oracle.action_probs(logits.detach(), masks).numpy(), g["probs"])
    np.testing.assert_array_equal(head_oracle.action_probs(logits.detach(), None).numpy(), g["probs_nomask"])
    assert np.array_equal(head_oracle.predict(logits.detach(), True, masks).numpy(), g["predict_det"])
    torch.manual_seed(4321)
    assert np.array_equal(head_oracle.predict(logits.detach(), False, masks).numpy(), g["predict_sample"])
    # sampled / argmax actions are always legal
    m = g["masks"][:, : logits.shape[1]]
    assert m[np.arange(m.shape[0]), g["sample_action"]].all()
    assert m[np.arange(m.shape[0]), g["det_action"]].all()


def test_jssp_batch_matches_reference(golden_dir):
    g = load(golden_dir, "jssp_batch.npz")
    ce, cemid = batch_oracle.conflicts_cliques(g["clq_mid"])
    assert np.array_equal(ce, g["clq_edges"]) and np.array_equal(cemid, g["clq_edges_mid"])
    for tag, conflicts, mie in (("att", "att", False), ("clique", "clique", False),
                                ("cliquemid", "clique", True)):
        b = batch_oracle.jssp_batch(g[f"{tag}.obs.features"], g[f"{tag}.obs.edge_index"],
                                    g[f"{tag}.obs.n_nodes"], g[f"{tag}.obs.n_edges"],
                                    conflicts=conflicts, mid_in_edges=mie)
        assert np.array_equal(b["src"], g[f"{tag}.src"]), tag
        assert np.array_equal(b["dst"], g[f"{tag}.dst"]), tag
        assert np.array_equal(b["type"], g[f"{tag}.type"]), tag
        assert np.array_equal(b["feat"], g[f"{tag}.feat"]), tag
        assert np.array_equal(b["batch_num_nodes"], g[f"{tag}.bnn"])
        assert np.array_equal(b["batch_num_edges"], g[f"{tag}.bne"])


@pytest.mark.parametrize("tag", ["clique", "cliquepad", "node", "att_max"])
def test_trunk_oracle_matches_reference_gnn_dgl(golden_dir, tag):
    """oracle.trunk_oracle.OracleGnnDGL + oracle.batch_oracle against the reference's own GnnDGL.forward
    (jssp/models/gnn_dgl.py:245-431 executed on the DGL shim): output and every parameter gradient."""
    from oracle import trunk_oracle
    g = load(golden_dir, "gnn_dgl.npz")
    n_j, n_m, B, Fdim, Hd, H, L, max_n_nodes, mie = (int(v) for v in g[f"{tag}.cfg"])
    conflicts, pooling = str(g[f"{tag}.conflicts"]), str(g[f"{tag}.pooling"])
    obs = {k: g[f"{tag}.obs.{k}"] for k in ("features", "edge_index", "n_nodes", "n_edges")}
    b = batch_oracle.jssp_batch(obs["features"], obs["edge_index"], obs["n_nodes"], obs["n_edges"],
                                conflicts=conflicts, mid_in_edges=bool(mie))
    b = batch_oracle.jssp_add_pool_and_machine_nodes(b, pooling, conflicts, n_m)
    P = {k[len(tag) + 3:]: T(g[k]) for k in g.files if k.startswith(tag + ".p.")}
    gnn = trunk_oracle.OracleGnnDGL(Fdim, pooling, max_n_nodes, n_m, 3, L, Hd, "gelu", H,
                                    n_edge_types=P["edge_embedder.0.weight"].shape[0], conflicts=conflicts)
    gnn.load_state_dict(P)
    NT = int(b["num_task_nodes"])
    ret = gnn.forward_cat(T(b["src"]), T(b["dst"]), T(b["type"]), T(b["feat"]), B, NT // B, NT)
    np.testing.assert_allclose(ret.detach().numpy(), g[f"{tag}.ret"], rtol=2e-5, atol=2e-6)
    (ret * T(g[f"{tag}.R"])).sum().backward()
    for k, p in gnn.named_parameters():
        ref = g[f"{tag}.g.{k}"]
        got = np.zeros_like(ref) if p.grad is None else p.grad.numpy()
        np.testing.assert_allclose(got, ref, rtol=2e-4, atol=2e-5 * max(1.0, np.abs(ref).max()), err_msg=k)


def test_gatv2_oracle_matches_graphconv_golden(golden_dir):
    """oracle.gat_oracle.gatv2_forward against psp/graph/graph_conv.py GraphConv(pyg=True) executed on the PyG
    shim (gatv2.npz).  The shim's GATv2Conv is itself a restatement of the third-party layer: this pins the
    adapter (constructor arguments, return conventions, parameter names) and keeps the two restatements honest
    against each other; the layer's parity stays "unpinned (third-party layer)"."""
    g = load(golden_dir, "gatv2.npz")
    P = {k[2:]: T(g[k]).requires_grad_(True) for k in g.files if k.startswith("p.")}
    assert set(P) == {"conv.att", "conv.bias", "conv.lin_l.weight", "conv.lin_l.bias", "conv.lin_r.weight",
                      "conv.lin_r.bias", "conv.lin_edge.weight"}
    x = T(g["x"]).requires_grad_(True)
    ea = T(g["ea"]).requires_grad_(True)
    ei = T(g["edge_index"])
    args_ = (P["conv.lin_l.weight"], P["conv.lin_l.bias"], P["conv.lin_r.weight"], P["conv.lin_r.bias"],
             P["conv.lin_edge.weight"], P["conv.att"], P["conv.bias"])
    out, alpha = gat_oracle.gatv2_forward(x, ei, ea, *args_)
    np.testing.assert_allclose(out.detach().numpy(), g["out"], rtol=1e-12, atol=1e-13)
    np.testing.assert_allclose(alpha.detach().numpy(), g["alpha"], rtol=1e-12, atol=1e-13)
    np.testing.assert_allclose(out[7].detach().numpy(), g["p.conv.bias"], rtol=0, atol=0)   # isolated target
    out_ne, _ = gat_oracle.gatv2_forward(x, ei, None, *args_)
    np.testing.assert_allclose(out_ne.detach().numpy(), g["out_noedge"], rtol=1e-12, atol=1e-13)
    names = list(P)
    grads = torch.autograd.grad((out * T(g["R"])).sum(), [P[k] for k in names] + [x, ea])
    for k, gr in zip(names, grads):
        np.testing.assert_allclose(gr.numpy(), g["g." + k], rtol=1e-10, atol=1e-12, err_msg=k)
    np.testing.assert_allclose(grads[-2].numpy(), g["gx"], rtol=1e-10, atol=1e-12)
    np.testing.assert_allclose(grads[-1].numpy(), g["gea"], rtol=1e-10, atol=1e-12)


PSP_TAGS = ["node272", "clique272", "nodeinv", "ragged", "clique_noself", "single"]


def psp_golden_case(g, tag):
    """(graphs, options, store order of the base edge types) of one psp_batch.npz case."""
    conflicts, pooling, _vnode, self_loops = (str(x) for x in g[f"{tag}.opts"])
    B = int(g[f"{tag}.n_graphs"])
    graphs = [dict(feat=g[f"{tag}.g{i}.feat"], resources=g[f"{tag}.g{i}.resources"],
                   res_cal=g[f"{tag}.g{i}.res_cal"], prec=g[f"{tag}.g{i}.prec"], rp=g[f"{tag}.g{i}.rp"],
                   rp_r=g[f"{tag}.g{i}.rp_r"]) for i in range(B)]
    return graphs, conflicts, pooling, self_loops == "1", [str(x) for x in g[f"{tag}.g0.stores_env"]]


@pytest.mark.parametrize("tag", PSP_TAGS)
def test_psp_batch_oracle_matches_reference_preprocess(golden_dir, tag):
    """oracle.psp_batch_oracle against psp/models/agent.py Agent.preprocess (+ rewire_internal, rewire,
    homogeneous_edges) executed on the stand-in container: every integer / attribute array bit for bit."""
    from oracle import psp_batch_oracle as pbo
    g = load(golden_dir, "psp_batch.npz")
    graphs, conflicts, pooling, self_loops, env_stores = psp_golden_case(g, tag)
    order = pbo.base_store_order(conflicts, env_stores)
    assert order == [str(x) for x in g[f"{tag}.g0.stores"]][:len(order)]
    out = pbo.psp_batch(graphs, conflicts, pooling, self_loops, store_order=order)
    for k in ("edge_index", "type", "rid", "att_rc", "att_rp", "res_cal_id"):
        ref = g[f"{tag}.out.{k}"]
        assert out[k].dtype == ref.dtype and out[k].shape == ref.shape, (k, out[k].dtype, ref.dtype)
        assert np.array_equal(out[k], ref), k
    assert out["batch_size"] == int(g[f"{tag}.out.batch_size"]) and out["num_nodes"] == int(g[f"{tag}.out.num_nodes"])
    assert np.array_equal(out["n_nodes"], g[f"{tag}.out.n_nodes"])
    assert np.array_equal(out["poolnodes"], g[f"{tag}.out.poolnodes"])
    assert np.array_equal(out["resource_nodes"], g[f"{tag}.out.resource_nodes"])
    ref_feat = g[f"{tag}.out.feat"]
    assert out["feat"].shape == ref_feat.shape
    keep = np.ones(ref_feat.shape[0], bool)
    keep[out["resource_nodes"]] = False                 # torch.empty rows in the reference (rewiring.py:132-141)
    assert np.array_equal(out["feat"][keep], ref_feat[keep])


@pytest.mark.parametrize("tag", ["node272", "clique272", "ragged_max", "single"])
def test_gnn_mp_oracle_matches_reference_trunk(golden_dir, tag):
    """oracle.trunk_oracle.OracleGnnMP (+ psp_batch_oracle) against the reference's GnnMP.forward / GnnFlat.forward
    executed on the shims (gnn_mp.npz): both outputs and every parameter gradient, reference state-dict keys."""
    from oracle import psp_batch_oracle as pbo
    from oracle import trunk_oracle
    g = load(golden_dir, "gnn_mp.npz")
    graphs, conflicts, pooling, self_loops, env_stores = psp_golden_case(g, tag)
    Hd, H, L, max_n_nodes, _max_res = (int(v) for v in g[f"{tag}.cfg"])
    b = pbo.psp_batch(graphs, conflicts, pooling, self_loops, store_order=pbo.base_store_order(conflicts, env_stores))
    gnn = trunk_oracle.OracleGnnMP(16, pooling, max_n_nodes, 3, L, Hd, "gelu", H, conflicts=conflicts)
    P = {k[len(tag) + 3:]: T(g[k]) for k in g.files if k.startswith(tag + ".p.")}
    assert sorted(P) == sorted(str(k) for k in g[f"{tag}.state_keys"])
    missing = gnn.load_state_dict(P, strict=False)
    assert not missing.missing_keys, missing
    n_nodes = b["n_nodes"].tolist()
    nf, ge = gnn(T(b["edge_index"]), T(b["type"]), T(b["att_rc"]), T(b["att_rp"]), T(b["feat"]), b["batch_size"],
                 b["num_nodes"], n_nodes[0] if len(n_nodes) == 1 else n_nodes, b["poolnodes"].tolist(),
                 b["resource_nodes"].tolist(), T(b["res_cal_id"]))
    np.testing.assert_allclose(nf.detach().numpy(), g[f"{tag}.node_features"], rtol=2e-5, atol=2e-6)
    np.testing.assert_allclose(ge.detach().numpy(), g[f"{tag}.graph_embedding"], rtol=2e-5, atol=2e-6)
    ((nf * T(g[f"{tag}.R1"])).sum() + (ge * T(g[f"{tag}.R2"])).sum()).backward()
    for k, p in gnn.named_parameters():
        ref = g[f"{tag}.g.{k}"]
        got = np.zeros_like(ref) if p.grad is None else p.grad.numpy()
        np.testing.assert_allclose(got, ref, rtol=3e-4, atol=3e-5 * max(1.0, np.abs(ref).max()), err_msg=k)


def test_ppo_oracle_matches_reference_lines(golden_dir):
    """oracle/ppo_oracle.py against ppo.npz = the reference's own source lines alg/ppo.py:276-299 (GAE) and
    :582-705 (minibatch loss) exec'ed by make_golden.py: bit-exact (same torch ops in the same order)."""
    from oracle import ppo_oracle as po
    g = np.load(os.path.join(golden_dir, "ppo.npz"))
    T = lambda a: torch.from_numpy(np.asarray(a))
    for tag in "abc":
        gamma, lam, sigma = (float(x) for x in g[f"gae.{tag}.cfg"])
        a, r = po.gae(T(g[f"gae.{tag}.rewards"]), T(g[f"gae.{tag}.values"]), T(g[f"gae.{tag}.dones"]),
                      T(g[f"gae.{tag}.next_value"]), T(g[f"gae.{tag}.next_done"]), gamma, lam, sigma)
        assert torch.equal(a, T(g[f"gae.{tag}.advantages"])) and torch.equal(r, T(g[f"gae.{tag}.returns"]))
    for tag in g["loss.tags"]:
        c = g[f"loss.{tag}.cfg"]
        lp, ent, v, logits = (T(g[f"loss.{tag}.{k}"]).requires_grad_(True) for k in ("logprob", "entropy", "value", "logits"))
        loss, st = po.ppo_loss(lp, ent, v, T(g[f"loss.{tag}.old_logprob"]), T(g[f"loss.{tag}.advantages"]).squeeze(1),
                               T(g[f"loss.{tag}.returns"]), T(g[f"loss.{tag}.old_values"]),
                               clip_coef=None if c[0] == "None" else float(c[0]), norm_adv=c[1] == "True",
                               clip_vloss=c[2] == "True", critic_loss=str(c[3]), ent_coef=float(c[4]),
                               vf_coef=float(c[5]), update=int(c[6]),
                               unmasked_probs=torch.distributions.Categorical(logits=logits).probs)
        assert torch.equal(loss.detach(), T(g[f"loss.{tag}.loss"])), tag
        for k in ("pg_loss", "v_loss", "entropy_loss", "approx_kl", "clipfrac"):
            assert torch.allclose(st[k].detach().float(), T(g[f"loss.{tag}.{k}"]).float(), rtol=0, atol=0), (tag, k)
        gs = torch.autograd.grad(loss, [lp, ent, v, logits], allow_unused=True)
        for got, like, k in zip(gs, (lp, ent, v, logits), ("g_logprob", "g_entropy", "g_value", "g_logits")):
            got = torch.zeros_like(like) if got is None else got
            assert torch.equal(got, T(g[f"loss.{tag}.{k}"])), (tag, k)
