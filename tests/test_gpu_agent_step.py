"""``agent.ppo_minibatch`` (wheatley_b200.agent_step: ONE autograd node, b200gat_agent_fwd / b200gat_agent_bwd) against
the module-by-module path it replaces - ``get_action_and_value`` (generic/agent.py:190-231) followed by
``wheatley_b200.ppo.ppo_loss`` (alg/ppo.py:582-705) - which the other GPU tests pin to the reference goldens and the
fp64 oracle.  Same kernels on the same operands: loss and statistics must agree to fp32 rounding (<= 1e-6 relative),
gradients to the bounds of tests/test_gpu_fused_trunk.py."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ATOMIC = ("attn", ".att", "fc_fij", "lin_edge", "edge_embedder", "bias", "edge_type", "rc_att", "rp_att")
KW = dict(clip_coef=0.2, norm_adv=True, critic_loss="l2", ent_coef=0.03, vf_coef=0.5)


@pytest.fixture(scope="module")
def wb():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import wheatley_b200
    return wheatley_b200


@pytest.fixture()
def tf32():
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    yield
    torch.backends.cuda.matmul.allow_tf32 = prev
    os.environ.pop("WHEATLEY_B200_TRUNK", None)


def rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max()) / max(float(b.abs().max()), 1e-6)


def targets(B, n, seed):
    g = torch.Generator().manual_seed(seed)
    masks = torch.rand(B, n, generator=g) < 0.4
    masks[:, 0] = True
    action = torch.multinomial(masks.float(), 1, generator=g).squeeze(1)
    return dict(masks=masks.cuda(), action=action.cuda(), adv=torch.randn(B, generator=g).cuda(),
                ret=torch.randn(B, generator=g).cuda(), old_lp=(-torch.rand(B, generator=g) * 2 - 3).cuda(),
                old_v=torch.randn(B, generator=g).cuda())


def compare(agent, make_obs, t, n_launch_fused_max=None):
    from wheatley_b200 import _lib
    from wheatley_b200.ppo import ppo_loss
    res = {}
    for mode in ("modules", "step"):
        for p in agent.parameters():
            p.grad = None
        obs = make_obs()
        n0 = _lib.launch_count()
        if mode == "modules":
            _, lp, ent, val, _ = agent.get_action_and_value(obs, t["action"], t["masks"])
            loss, stats = ppo_loss(lp, ent, val, t["old_lp"], t["adv"], t["ret"], t["old_v"], **KW)
        else:
            loss, stats = agent.ppo_minibatch(obs, t["action"], t["masks"], t["old_lp"], t["adv"], t["ret"],
                                              t["old_v"], **KW)
        (loss * 1.25).backward()                  # an upstream gradient != 1
        torch.cuda.synchronize()
        res[mode] = (loss.detach().clone(), stats.detach().clone(),
                     {k: p.grad.detach().clone() for k, p in agent.named_parameters() if p.grad is not None},
                     _lib.launch_count() - n0)
    assert rel(res["step"][0], res["modules"][0]) < 1e-6
    assert rel(res["step"][1], res["modules"][1]) < 1e-5
    assert set(res["step"][2]) == set(res["modules"][2])
    for k, g in res["step"][2].items():
        tol = 3e-4 if any(s in k for s in ATOMIC) else 1e-4
        assert rel(g, res["modules"][2][k]) < tol, k
    return res


@pytest.mark.parametrize("conflicts,pooling", [("clique", "learn"), ("node", "learn"), ("clique", "max")])
def test_jssp_agent_ppo_minibatch_equals_module_path(wb, tf32, conflicts, pooling):
    from wheatley_b200.agent import JsspAgent
    from wheatley_b200.batch import jssp_batch_from_observation
    from wheatley_b200.workload import jssp_observation_batch
    n_j, n_m, B, L = 12, 5, 3, 3
    obs = jssp_observation_batch(n_j, n_m, B, seed0=4, progress=0.4)
    torch.manual_seed(2)
    agent = JsspAgent(obs["features"].shape[2], n_j * n_m, n_m, 64, L, 4, graph_pooling=pooling,
                      conflicts=conflicts).cuda()
    mie = conflicts == "node"
    if mie:                          # machine-node edge types need the wider table
        from wheatley_b200.gnn_dgl import GnnDGL
        agent.gnn = GnnDGL(obs["features"].shape[2], "gatv2", pooling, False, n_j * n_m, n_m, 3, L, 64, "gelu", 4,
                           conflicts="node", mid_in_edges=True, split_output=True).cuda()

    def make_obs():
        jb, _ = jssp_batch_from_observation(obs, "cuda", n_j * n_m, conflicts=conflicts, graph_pooling=pooling,
                                            mid_in_edges=mie, max_n_machines=n_m if mie else 0)
        return jb

    res = compare(agent, make_obs, targets(B, n_j * n_m, 1))
    assert res["step"][3] <= res["modules"][3] + 8      # the same kernels + a handful of glue launches


@pytest.mark.parametrize("conflicts", ["node", "clique"])
def test_psp_agent_ppo_minibatch_equals_module_path(wb, tf32, conflicts):
    from wheatley_b200.agent import PspAgent
    from wheatley_b200.wire import PspArena
    from wheatley_b200.workload import psp_taillard_graph
    B = 7
    graphs = [psp_taillard_graph(4, 4, 3 + b, n_rp=b % 3, progress=0.1 * b) for b in range(B)]
    torch.manual_seed(0)
    agent = PspAgent(18, 4, hidden_dim=32, n_layers=4, n_heads=4, conflicts=conflicts).cuda()
    arena = PspArena.from_graphs(graphs)
    compare(agent, lambda: agent.preprocess(arena), targets(B, 18, 2))


def test_ppo_minibatch_without_tf32_takes_the_module_path(wb):
    """fp32-exact matmuls (``allow_tf32=False``): the fused step does not apply; same API, module path."""
    from wheatley_b200.agent import JsspAgent
    from wheatley_b200.batch import jssp_batch_from_observation
    from wheatley_b200.workload import jssp_observation_batch
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        obs = jssp_observation_batch(8, 3, 2, seed0=0)
        torch.manual_seed(0)
        agent = JsspAgent(obs["features"].shape[2], 24, 3, 64, 2, 4).cuda()
        jb, _ = jssp_batch_from_observation(obs, "cuda", 24, conflicts="clique", graph_pooling="learn")
        t = targets(2, 24, 3)
        loss, stats = agent.ppo_minibatch(jb, t["action"], t["masks"], t["old_lp"], t["adv"], t["ret"], **KW)
        loss.backward()
        assert torch.isfinite(loss) and stats.shape == (8,)
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
