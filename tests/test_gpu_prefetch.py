"""BatchPrefetcher / deferred plan build: the pipelined batches must be the same plans, and give the same losses
(to the bit) and gradients, as batches built directly on the current stream."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def wb():
    import wheatley_b200
    return wheatley_b200


def _host_batch(n_j, n_m, B, seed):
    from wheatley_b200.workload import jssp_batch
    hb = jssp_batch(n_j, n_m, B, seed0=seed, progress=0.25)
    g = torch.Generator().manual_seed(seed)
    masks = torch.rand(B, n_j * n_m, generator=g) < 0.4
    masks[:, 0] = True
    action = torch.multinomial(masks.float(), 1, generator=g).squeeze(1)
    host = dict(src=hb["src"], dst=hb["dst"], etype=hb["etype"], feat=hb["feat"], masks=masks, action=action,
                adv=torch.randn(B, generator=g), ret=torch.randn(B, generator=g))
    return {k: v.contiguous().pin_memory() for k, v in host.items()}, hb["n_nodes"], hb["n_task_nodes"]


def _step(agent, plan, d, B, n, NT):
    from wheatley_b200.gnn_dgl import JsspBatch
    from wheatley_b200.ppo import ppo_like_loss
    jb = JsspBatch(plan, d["etype"], d["feat"], B, n, NT, poolnodes=torch.arange(NT, NT + B, device="cuda"))
    for p in agent.parameters():
        p.grad = None
    _, lp, ent, val, _ = agent.get_action_and_value(jb, d["action"], d["masks"])
    loss = ppo_like_loss(lp, ent, val, d["adv"], d["ret"])
    loss.backward()
    return loss.detach().clone(), [p.grad.detach().clone() for p in agent.parameters() if p.grad is not None]


def test_deferred_prepare_builds_the_same_view(wb):
    host, N, _ = _host_batch(10, 5, 3, seed=2)
    d = {k: v.cuda() for k, v in host.items()}
    a = wb.GraphPlan(d["src"], d["dst"], N).prepare(d["etype"])
    b = wb.GraphPlan(d["src"], d["dst"], N)
    b.prepare(d["etype"], defer=True)
    vb = next(iter(b._clique_views.values()))
    assert vb._pending is not None and vb.n_blocks == 0          # queued, counts not read yet
    b.finish()
    va = next(iter(a._clique_views.values()))
    assert vb._pending is None and va.n_blocks == vb.n_blocks > 0
    assert (va.max_block, va.n_res_edges, va.n_members, va.s_elems) == \
           (vb.max_block, vb.n_res_edges, vb.n_members, vb.s_elems)
    assert a.stats_host() == b.stats_host()
    for name in ("node_blk", "node_run", "node_rank"):
        assert torch.equal(getattr(va, name), getattr(vb, name)), name
    nb = va.n_blocks                          # blk_type is allocated for the upper bound N: the tail is unwritten
    assert torch.equal(va.blk_type[:nb], vb.blk_type[:nb])
    for x, y in zip(va.r, vb.r):
        assert torch.equal(x, y)
    assert len(b.tensors()) >= 20


@pytest.mark.parametrize("side_stream", [False, True])
def test_prefetched_batches_train_identically(wb, side_stream):
    """Four different minibatches through the pipeline (next batch queued while the current one trains) against
    the same four built directly, on the bench's dense path (tcgen05 TF32): the forward is deterministic, so the
    losses must agree to the bit; the backward accumulates a few small gradients with atomics (order-dependent
    rounding; 7e-5 of a 3e-6-sized gradient was seen), so gradients are compared to 1e-3 of their scale."""
    from wheatley_b200.agent import JsspAgent
    from wheatley_b200.prefetch import BatchPrefetcher
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        n_j, n_m, B = 15, 8, 4
        batches = [_host_batch(n_j, n_m, B, seed=s) for s in range(4)]
        torch.manual_seed(0)
        agent = JsspAgent(batches[0][0]["feat"].shape[1], n_j * n_m, n_m, 64, 3, 4).cuda()
        direct = []
        for host, N, NT in batches:
            d = {k: v.cuda() for k, v in host.items()}
            direct.append(_step(agent, wb.GraphPlan(d["src"], d["dst"], N), d, B, n_j * n_m, NT))
        torch.cuda.synchronize()
        pf = BatchPrefetcher("cuda:0", side_stream=side_stream)
        pf.submit(batches[0][0], batches[0][1])
        piped = []
        for i, (host, N, NT) in enumerate(batches):
            d, plan = pf.get()
            if i + 1 < len(batches):
                pf.submit(batches[i + 1][0], batches[i + 1][1])
            piped.append(_step(agent, plan, d, B, n_j * n_m, NT))    # no sync: the next get() overlaps with this
        torch.cuda.synchronize()
        for (l0, g0), (l1, g1) in zip(direct, piped):
            assert torch.equal(l0, l1)
            assert len(g0) == len(g1)
            for a, b in zip(g0, g1):
                assert float((a - b).abs().max()) <= 1e-3 * max(float(a.abs().max()), 1e-6)
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old


def test_prefetcher_get_without_submit_raises(wb):
    from wheatley_b200.prefetch import BatchPrefetcher
    with pytest.raises(wb.B200GatError):
        BatchPrefetcher("cuda:0").get()


def test_flat_adam_on_the_agent_matches_torch_adam(wb):
    """ppo.FlatAdam (fused Adam on one flat parameter) against torch.optim.Adam(fused) on the whole agent,
    three training steps on the same batch and the same gradients: same parameters."""
    from wheatley_b200.agent import JsspAgent
    from wheatley_b200.ppo import FlatAdam
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        n_j, n_m, B = 10, 5, 3
        host, N, NT = _host_batch(n_j, n_m, B, seed=1)
        d = {k: v.cuda() for k, v in host.items()}
        torch.manual_seed(0)
        a = JsspAgent(host["feat"].shape[1], n_j * n_m, n_m, 64, 2, 4).cuda()
        b = JsspAgent(host["feat"].shape[1], n_j * n_m, n_m, 64, 2, 4).cuda()
        b.load_state_dict(a.state_dict())
        oa = FlatAdam(a.parameters(), lr=1e-3)
        ob = torch.optim.Adam(b.parameters(), lr=1e-3, fused=True, capturable=True)
        plan = wb.GraphPlan(d["src"], d["dst"], N)
        for it in range(3):
            _step(a, plan, d, B, n_j * n_m, NT)
            # the same gradients for both optimizers (Adam turns the rounding noise of a theoretically zero
            # gradient into +-lr steps, so two separately computed backward passes may not be compared)
            for pa, pb in zip(a.parameters(), b.parameters()):
                pb.grad = None if pa.grad is None else pa.grad.clone()
            oa.step()
            ob.step()
            for (k, p), q in zip(a.named_parameters(), b.parameters()):
                assert float((p - q).abs().max()) <= 1e-6 * max(float(q.abs().max()), 1e-3), (it, k)
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old


def test_psp_batches_prefetched_on_the_side_stream_train_identically(wb):
    """PSP arenas through ``submit_psp`` on the side stream, two batches in flight, graphs whose pool node is a hub
    (> 256 in-edges: the split-hub path with its zero-filled arrival counters, which ``plan.finish()`` queues on the
    side stream after the event recorded at submit time): the losses equal those of batches built directly on the
    main stream, bit for bit (the forward is deterministic), step after step."""
    from wheatley_b200.agent import PspAgent
    from wheatley_b200.prefetch import BatchPrefetcher
    from wheatley_b200.wire import PspArena
    from wheatley_b200.workload import psp_taillard_graph
    n_j, n_m, B, steps = 20, 15, 2, 6
    arenas = [PspArena.from_graphs([psp_taillard_graph(n_j, n_m, 10 * s + b, progress=0.0) for b in range(B)])
              for s in range(steps)]
    max_n = n_j * n_m + 2
    torch.manual_seed(0)
    agent = PspAgent(max_n, n_m, hidden_dim=64, n_layers=2, conflicts="node").cuda()
    g = torch.Generator().manual_seed(3)
    masks = torch.rand(B, max_n, generator=g) < 0.3
    masks[:, 0] = True
    extras = dict(masks=masks, action=torch.multinomial(masks.float(), 1, generator=g).squeeze(1),
                  adv=torch.randn(B, generator=g), ret=torch.randn(B, generator=g),
                  old_lp=-torch.rand(B, generator=g) - 4.0, old_v=torch.randn(B, generator=g))
    extras = {k: v.contiguous().pin_memory() for k, v in extras.items()}
    kw = dict(clip_coef=0.2, norm_adv=True, critic_loss="l2", ent_coef=0.05, vf_coef=0.5)

    def step(obs, d):
        for p in agent.parameters():
            p.grad = None
        loss, _ = agent.ppo_minibatch(obs, d["action"], d["masks"], d["old_lp"], d["adv"], d["ret"], d["old_v"], **kw)
        loss.backward()
        return loss.detach().clone()

    dd = {k: v.cuda() for k, v in extras.items()}
    direct = [step(agent.preprocess(a), dd) for a in arenas]
    torch.cuda.synchronize()
    assert direct[0].isfinite() and agent.preprocess(arenas[0])[0].plan.stats_host()[1] > 256       # a hub
    pf = BatchPrefetcher("cuda:0", side_stream=True)
    pf.submit_psp(agent, arenas[0], extras)
    pf.submit_psp(agent, arenas[1], extras)
    piped = []
    for i in range(steps):
        d, _ = pf.get()
        if i + 2 < steps:
            pf.submit_psp(agent, arenas[i + 2], extras)
        piped.append(step(d["batch"], d))                   # no sync between the steps
    torch.cuda.synchronize()
    for a, b in zip(direct, piped):
        assert torch.equal(a, b), (float(a), float(b))
