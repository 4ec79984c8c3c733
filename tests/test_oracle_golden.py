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
