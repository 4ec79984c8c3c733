"""GPU parity for the integer paths (graph plan: bit-exact) and the masked action head
(golden vectors produced by the reference's own Agent methods; fp32)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def wb():
    import wheatley_b200
    return wheatley_b200


def check_plan(wb, src, dst, n):
    plan = wb.GraphPlan(src.cuda(), dst.cuda(), n)
    s, d = src.numpy(), dst.numpy()
    E = s.shape[0]
    order_in = np.argsort(d, kind="stable")
    order_out = np.argsort(s, kind="stable")
    in_ptr = np.concatenate([[0], np.cumsum(np.bincount(d, minlength=n))])
    out_ptr = np.concatenate([[0], np.cumsum(np.bincount(s, minlength=n))])
    inv = np.empty(E, dtype=np.int64)
    inv[order_in] = np.arange(E)
    assert np.array_equal(plan.in_ptr.cpu().numpy(), in_ptr)
    assert np.array_equal(plan.out_ptr.cpu().numpy(), out_ptr)
    assert np.array_equal(plan.in_eid.cpu().numpy(), order_in)
    assert np.array_equal(plan.in_src.cpu().numpy(), s[order_in])
    assert np.array_equal(plan.out_dst.cpu().numpy(), d[order_out])
    assert np.array_equal(plan.out_pos.cpu().numpy(), inv[order_out])
    zero, max_in, bad, max_out = plan.stats_host()
    assert zero == int((np.bincount(d, minlength=n) == 0).sum())
    assert max_in == (int(np.bincount(d, minlength=n).max()) if E else 0)
    assert max_out == (int(np.bincount(s, minlength=n).max()) if E else 0)
    assert bad == 0
    for order in (plan.dst_order, plan.src_order):           # locality orders are permutations
        assert np.array_equal(np.sort(order.cpu().numpy()), np.arange(n))
    return plan


@pytest.mark.parametrize("n,e,hub", [(1, 1, None), (5, 0, None), (37, 500, None), (1000, 30000, 7),
                                     (5000, 200001, 0), (3, 4099, 1)])
def test_plan_is_stable_counting_sort(wb, n, e, hub):
    rng = np.random.default_rng(n + e)
    src = rng.integers(0, n, e)
    dst = rng.integers(0, n, e)
    if hub is not None and e:
        dst[: e // 2] = hub          # in-degree far above the warp-sort limit
        src[e // 3:] = hub
    check_plan(wb, torch.from_numpy(src).long(), torch.from_numpy(dst).long(), n)


def test_plan_taillard_100x20(wb):
    from wheatley_b200.workload import jssp_batch
    b = jssp_batch(100, 20, 2, progress=0.5)
    plan = check_plan(wb, b["src"], b["dst"], b["n_nodes"])
    et = plan.sort_edge_ints(b["etype"].cuda())
    assert np.array_equal(et.cpu().numpy(), b["etype"].numpy()[np.argsort(b["dst"].numpy(), kind="stable")])
    rows = torch.randn(b["n_edges"], 5)
    got = plan.sort_edge_rows(rows.cuda())
    assert torch.equal(got.cpu(), rows[torch.from_numpy(np.argsort(b["dst"].numpy(), kind="stable"))])


def test_plan_counts_out_of_range_edges(wb):
    plan = wb.GraphPlan(torch.tensor([0, 9, 1]).cuda(), torch.tensor([1, 1, -1]).cuda(), 3)
    assert plan.stats_host()[2] == 2
    with pytest.raises(wb.B200GatError):
        plan.validate()


def test_head_golden(wb, golden_dir):
    g = np.load(os.path.join(golden_dir, "head.npz"))
    T = lambda k: torch.from_numpy(g[k])
    from wheatley_b200 import agent_head
    nf = T("node_features").cuda().requires_grad_(True)
    logits = (nf @ T("w_action").cuda().t() + T("b_action").cuda()).squeeze(-1)
    masks = T("masks").cuda()
    a, lp, ent = agent_head.action_logprob_entropy(logits, None, masks, deterministic=True)
    assert torch.equal(a.cpu(), T("det_action"))
    np.testing.assert_allclose(lp.detach().cpu().numpy(), g["det_logprob"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(ent.detach().cpu().numpy(), g["det_entropy"], rtol=1e-5, atol=1e-6)
    _, lp2, ent2 = agent_head.action_logprob_entropy(logits, T("action_in").cuda(), masks)
    np.testing.assert_allclose(lp2.detach().cpu().numpy(), g["eval_logprob"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(ent2.detach().cpu().numpy(), g["eval_entropy"], rtol=1e-5, atol=1e-6)
    ((lp2 * T("wl").cuda()).sum() + (ent2 * T("we").cuda()).sum()).backward()
    np.testing.assert_allclose(nf.grad.cpu().numpy(), g["g_nf"], rtol=1e-4, atol=1e-6)
    _, lp3, ent3 = agent_head.action_logprob_entropy(logits.detach(), T("action_in").cuda(), None)
    np.testing.assert_allclose(lp3.cpu().numpy(), g["nomask_logprob"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(ent3.cpu().numpy(), g["nomask_entropy"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(agent_head.action_probs(logits.detach(), masks).cpu().numpy(), g["probs"],
                               rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(agent_head.action_probs(logits.detach(), None).cpu().numpy(),
                               g["probs_nomask"], rtol=1e-5, atol=1e-7)
    assert torch.equal(agent_head.predict_action(logits, True, masks).cpu(), T("predict_det"))
    # sampled actions are always legal and their log-prob is the gathered one
    torch.manual_seed(0)
    a4, lp4, _ = agent_head.action_logprob_entropy(logits.detach(), None, masks, deterministic=False)
    m = g["masks"][:, : logits.shape[1]]
    assert m[np.arange(m.shape[0]), a4.cpu().numpy()].all()
    a5 = agent_head.predict_action(logits, False, masks)
    assert m[np.arange(m.shape[0]), a5.cpu().numpy()].all()


@pytest.mark.parametrize("B,n", [(1, 1), (7, 33), (64, 2000), (3, 4097)])
def test_head_matches_oracle(wb, B, n):
    from oracle import head_oracle
    from wheatley_b200 import agent_head
    g = torch.Generator().manual_seed(B * n)
    logits = 3 * torch.randn(B, n, generator=g)
    masks = torch.rand(B, n + 5, generator=g) < 0.3           # wider than the logits: cropped (:202-203)
    masks[:, 0] = True
    action = torch.multinomial(masks[:, :n].float(), 1, generator=g).squeeze(1)
    wl, we = torch.randn(B, generator=g), torch.randn(B, generator=g)
    lr = logits.double().requires_grad_(True)
    _, lp_r, ent_r, _ = head_oracle.action_and_logprob(lr, action, masks)
    ((lp_r * wl).sum() + (ent_r * we).sum()).backward()
    lg = logits.cuda().requires_grad_(True)
    _, lp, ent = agent_head.action_logprob_entropy(lg, action.cuda(), masks.cuda())
    ((lp * wl.cuda()).sum() + (ent * we.cuda()).sum()).backward()
    np.testing.assert_allclose(lp.detach().cpu().numpy(), lp_r.detach().numpy(), rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(ent.detach().cpu().numpy(), ent_r.detach().numpy(), rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(lg.grad.cpu().numpy(), lr.grad.numpy(), rtol=1e-4, atol=1e-6)
    a_det, _, _ = agent_head.action_logprob_entropy(logits.cuda(), None, masks.cuda(), deterministic=True)
    a_ref, _, _, _ = head_oracle.action_and_logprob(logits, None, masks, deterministic=True)
    assert torch.equal(a_det.cpu(), a_ref)
    # generic log-softmax backward (logp used by arbitrary downstream code)
    lg2 = logits.cuda().requires_grad_(True)
    logp, _, _ = wb.masked_log_softmax(lg2, masks.cuda())
    W = torch.randn(B, n, generator=g)
    (logp * W.cuda() * masks[:, :n].cuda()).sum().backward()
    lr2 = logits.double().requires_grad_(True)
    ml = torch.where(masks[:, :n], lr2, torch.tensor(-1e20, dtype=torch.float64))
    ((ml - ml.logsumexp(-1, keepdim=True)) * W.double() * masks[:, :n]).sum().backward()
    np.testing.assert_allclose(lg2.grad.cpu().numpy(), lr2.grad.numpy(), rtol=1e-4, atol=1e-5)


def test_trunk_and_agent_match_oracle(wb):
    """Whole hot path: 3-graph batch through GnnDGL (3 layers) + heads + loss, fwd and bwd."""
    from oracle import trunk_oracle
    from wheatley_b200.agent import JsspAgent
    from wheatley_b200.gnn_dgl import JsspBatch
    from wheatley_b200.ppo import ppo_like_loss
    from wheatley_b200.workload import jssp_batch
    n_j, n_m, B, L, Hd, H = 8, 5, 4, 3, 32, 4
    hb = jssp_batch(n_j, n_m, B, seed0=3, progress=0.3)
    F = hb["feat"].shape[1]
    torch.manual_seed(1)
    agent = JsspAgent(F, n_j * n_m + 3, n_m, Hd, L, H).cuda()          # max_n_nodes > n_nodes: padded
    gnn = trunk_oracle.OracleGnnDGL(F, "learn", n_j * n_m + 3, n_m, 3, L, Hd, "gelu", H)
    ref = trunk_oracle.OracleAgent(gnn, F + Hd * (L + 1)).double()
    ref.load_state_dict({k: v.double().cpu() for k, v in agent.state_dict().items()})
    g = torch.Generator().manual_seed(1)
    masks = torch.rand(B, n_j * n_m + 3, generator=g) < 0.4
    masks[:, 0] = True
    masks[:, n_j * n_m:] = False
    action = torch.multinomial(masks.float(), 1, generator=g).squeeze(1)
    adv, ret = torch.randn(B, generator=g), torch.randn(B, generator=g)
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        plan = wb.GraphPlan(hb["src"].cuda(), hb["dst"].cuda(), hb["n_nodes"])
        NT = hb["n_task_nodes"]
        jb = JsspBatch(plan, hb["etype"].cuda(), hb["feat"].cuda(), B, n_j * n_m, NT,
                       poolnodes=torch.arange(NT, NT + B).cuda())
        _, lp, ent, val, unmasked = agent.get_action_and_value(jb, action.cuda(), masks.cuda())
        loss = ppo_like_loss(lp, ent, val, adv.cuda(), ret.cuda())
        loss.backward()
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
    args_ = (hb["src"], hb["dst"], hb["etype"], hb["feat"].double(), B, n_j * n_m, NT)
    lp_r, ent_r, val_r = ref.evaluate(args_, action, masks)
    loss_r = trunk_oracle.ppo_like_loss(lp_r, ent_r, val_r, adv.double(), ret.double())
    loss_r.backward()

    def rel(a, b):
        return float((a.detach().double().cpu() - b.detach()).abs().max()) / max(float(b.abs().max()), 1e-3)
    assert rel(lp, lp_r) < 1e-4 and rel(ent, ent_r) < 1e-4 and rel(val, val_r) < 1e-4
    assert rel(loss, loss_r) < 1e-4
    refp = dict(ref.named_parameters())
    for k, p in agent.named_parameters():
        if refp[k].grad is None:
            continue
        assert rel(p.grad, refp[k].grad) < 2e-4, k
    # checkpoint layout round trip (generic/agent.py:158-172)
    d = agent.save_dict()
    assert set(d) == {"gnn", "value_net", "action_net"}
    agent.load_dict(d)
