"""PPO update tail on the device: b200gat_gae and b200gat_ppo_loss against ppo.npz - the reference's
own source lines alg/ppo.py:276-299 and :582-705 executed by tests/golden/make_golden.py - and against the oracle
restatement on larger seeded inputs.  GAE is integer-like bookkeeping in fp32 with a fixed operation order:
bit-exact.  Loss, statistics and gradients: 1e-5 relative (fp32 sums in a different order)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def wb():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import wheatley_b200
    return wheatley_b200


def T(a):
    return torch.from_numpy(np.asarray(a))


def rel(a, b):
    a, b = a.detach().double().cpu().reshape(-1), torch.as_tensor(b).detach().double().cpu().reshape(-1)
    return float((a - b).abs().max()) / max(float(b.abs().max()), 1e-6)


def test_gae_bit_exact_against_reference_lines(wb, golden_dir):
    from wheatley_b200.ppo import gae
    g = np.load(os.path.join(golden_dir, "ppo.npz"))
    for tag in "abc":
        gamma, lam, sigma = (float(x) for x in g[f"gae.{tag}.cfg"])
        a, r = gae(*(T(g[f"gae.{tag}.{k}"]).cuda() for k in ("rewards", "values", "dones", "next_value", "next_done")),
                   gamma, lam, sigma)
        assert torch.equal(a.cpu(), T(g[f"gae.{tag}.advantages"])), tag
        assert torch.equal(r.cpu(), T(g[f"gae.{tag}.returns"])), tag


def test_gae_rollout_size_against_oracle(wb):
    from oracle import ppo_oracle as po
    from wheatley_b200.ppo import gae
    gen = torch.Generator().manual_seed(5)
    Tn, E, R = 1800, 10, 1                           # 272.sh: --n_steps_episode 1800 --n_workers 10
    rewards, values = torch.randn(Tn, E, R, generator=gen), torch.randn(Tn, E, R, generator=gen)
    dones = (torch.rand(Tn, E, generator=gen) < 0.05).float()
    nv, nd = torch.randn(E, R, generator=gen), (torch.rand(E, generator=gen) < 0.5).float()
    a_r, r_r = po.gae(rewards, values, dones, nv, nd, 0.995, 0.95, 1.3)
    a, r = gae(rewards.cuda(), values.cuda(), dones.cuda(), nv.cuda(), nd.cuda(), 0.995, 0.95, 1.3)
    assert torch.equal(a.cpu(), a_r) and torch.equal(r.cpu(), r_r)


def _cfg(c):
    return dict(clip_coef=None if c[0] == "None" else float(c[0]), norm_adv=c[1] == "True", clip_vloss=c[2] == "True",
                critic_loss=str(c[3]), ent_coef=float(c[4]), vf_coef=float(c[5]), update=int(c[6]))


def test_ppo_loss_against_reference_lines(wb, golden_dir):
    from wheatley_b200.ppo import STATS, ppo_loss
    g = np.load(os.path.join(golden_dir, "ppo.npz"))
    for tag in g["loss.tags"]:
        cfg = _cfg(g[f"loss.{tag}.cfg"])
        lp, ent, v, logits = (T(g[f"loss.{tag}.{k}"]).cuda().requires_grad_(True)
                              for k in ("logprob", "entropy", "value", "logits"))
        probs = torch.distributions.Categorical(logits=logits).probs
        loss, stats = ppo_loss(lp, ent, v, T(g[f"loss.{tag}.old_logprob"]).cuda(),
                               T(g[f"loss.{tag}.advantages"]).cuda().squeeze(1), T(g[f"loss.{tag}.returns"]).cuda(),
                               T(g[f"loss.{tag}.old_values"]).cuda(), unmasked_probs=probs, **cfg)
        assert rel(loss, g[f"loss.{tag}.loss"]) < 1e-5, tag
        st = dict(zip(STATS, stats.cpu().tolist()))
        for k in ("pg_loss", "v_loss", "entropy_loss", "approx_kl", "clipfrac"):
            assert abs(st[k] - float(g[f"loss.{tag}.{k}"])) <= 1e-5 * max(1.0, abs(float(g[f"loss.{tag}.{k}"]))), (tag, k)
        gs = torch.autograd.grad(loss * 1.5, [lp, ent, v, logits], allow_unused=True)      # upstream scale != 1
        for got, like, k in zip(gs, (lp, ent, v, logits), ("g_logprob", "g_entropy", "g_value", "g_logits")):
            got = torch.zeros_like(like) if got is None else got
            ref = 1.5 * T(g[f"loss.{tag}.{k}"])
            assert float((got.cpu() - ref).abs().max()) <= 1e-5 * max(float(ref.abs().max()), 1e-3), (tag, k)


@pytest.mark.parametrize("B", [128, 1536, 5000])
def test_ppo_loss_minibatch_sizes_against_oracle(wb, B):
    from oracle import ppo_oracle as po
    from wheatley_b200.ppo import ppo_loss
    gen = torch.Generator().manual_seed(B)
    lp = (-torch.rand(B, generator=gen) * 3).requires_grad_(True)
    old = lp.detach() + 0.3 * torch.randn(B, generator=gen)
    ent, v = torch.rand(B, generator=gen).requires_grad_(True), torch.randn(B, 1, generator=gen).requires_grad_(True)
    adv, ret = torch.randn(B, generator=gen) * 3 + 1, torch.randn(B, 1, generator=gen)
    kw = dict(clip_coef=0.2, norm_adv=True, critic_loss="l2", ent_coef=0.01, vf_coef=0.5)
    loss_r, st_r = po.ppo_loss(lp, ent, v, old, adv, ret, **kw)
    g_r = torch.autograd.grad(loss_r, [lp, ent, v])
    d = [t.detach().cuda().requires_grad_(True) for t in (lp, ent, v)]
    loss, stats = ppo_loss(d[0], d[1], d[2], old.cuda(), adv.cuda(), ret.cuda(), **kw)
    assert rel(loss, loss_r) < 1e-5
    assert abs(float(stats[6]) - float(st_r["adv_mean"])) < 1e-5 and rel(stats[7], st_r["adv_std"]) < 1e-5
    for got, ref in zip(torch.autograd.grad(loss, d), g_r):
        assert rel(got, ref) < 1e-5


def test_sharded_loss_shares_sum_to_the_single_process_loss(wb):
    """a minibatch split into UNEQUAL shards, every shard evaluated with the all-reduced advantage moments
    (here: summed by hand, what dist.all_reduce(SUM) does) - losses, statistics and gradients of the shards add up
    to the single-process values, so the gradient exchange is a plain SUM."""
    from wheatley_b200 import _lib
    from wheatley_b200.ppo import _PpoLoss, adv_moments, ppo_loss
    gen = torch.Generator().manual_seed(3)
    B = 100
    lp = (-torch.rand(B, generator=gen)).cuda()
    old, ent, v = lp + 0.2 * torch.randn(B, generator=gen).cuda(), torch.rand(B, generator=gen).cuda(), torch.randn(B, generator=gen).cuda()
    adv, ret = torch.randn(B, generator=gen).cuda() * 2, torch.randn(B, generator=gen).cuda()
    kw = dict(clip_coef=0.2, norm_adv=True, clip_vloss=False, critic_loss="l2", ent_coef=0.02, vf_coef=0.5)
    full = [t.clone().requires_grad_(True) for t in (lp, ent, v)]
    loss, stats = ppo_loss(full[0], full[1], full[2], old, adv, ret, **kw)
    g_full = torch.autograd.grad(loss, full)
    bounds = [(0, 13), (13, 70), (70, 100)]
    moments = sum(adv_moments(adv[a:b]) for a, b in bounds)
    assert float(moments[0]) == B
    tot_loss, tot_stats, g_parts = 0.0, 0.0, []
    for a, b in bounds:
        part = [t[a:b].clone().requires_grad_(True) for t in (lp, ent, v)]
        l, s = _PpoLoss.apply(part[0], part[1], part[2], old[a:b], adv[a:b], ret[a:b], None, moments, kw)
        tot_loss, tot_stats = tot_loss + l.detach(), tot_stats + s[:6]
        g_parts.append(torch.autograd.grad(l, part))
    assert rel(tot_loss, loss) < 1e-5 and rel(tot_stats, stats[:6]) < 1e-5
    for k in range(3):
        assert rel(torch.cat([g[k] for g in g_parts]), g_full[k]) < 1e-5
