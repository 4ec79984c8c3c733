"""TEST INFRASTRUCTURE (oracle) - CPU restatement of the PPO update tail of the reference.

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this.  Plain PyTorch, fp32 like the
reference; pinned by tests/golden/ppo.npz, which tests/golden/make_golden.py produces by EXECUTING the reference's
own source lines (alg/ppo.py:276-299 and :582-705 are in the middle of PPO.collect_rollouts / PPO.train, and the
module needs gymnasium, so the lines are read from /root/reference at generation time and exec'ed in a namespace
that supplies the variables those methods have in scope)."""
import torch


def gae(rewards, values, dones, next_value, next_done, gamma, gae_lambda, sigma=1.0):
    """alg/ppo.py:276-299.  rewards / values [T, E, R], dones [T, E], next_value [E, R], next_done [E]."""
    T = rewards.shape[0]
    advantages = torch.empty_like(rewards)
    lastgaelam = 0
    for t in reversed(range(T)):                                                  # :279
        if t == T - 1:                                                            # :280-282
            nextnonterminal = 1.0 - next_done
            nextvalues = next_value
        else:                                                                     # :283-285
            nextnonterminal = 1.0 - dones[t + 1]
            nextvalues = values[t + 1]
        delta = (rewards[t] + gamma * nextvalues * nextnonterminal.unsqueeze(1) - values[t]) / sigma    # :287-291
        advantages[t] = lastgaelam = delta + gamma * gae_lambda * nextnonterminal.unsqueeze(1) * lastgaelam  # :292-298
    return advantages, advantages + values                                        # :299


def ppo_loss(newlogprob, entropy, newvalue, old_logprobs, advantages, returns, old_values=None, *, clip_coef=0.2,
             norm_adv=True, clip_vloss=False, critic_loss="l2", ent_coef=0.0, vf_coef=0.5, update=1,
             unmasked_probs=None):
    """alg/ppo.py:582-705 with ``reward_dim == 1`` advantages already aggregated, no two_hot / hl_gauss / symlog
    (the defaults).  Returns ``(loss, dict of the logged scalars)``."""
    logratio = newlogprob - old_logprobs                                          # :582
    ratio = logratio.exp()                                                        # :583
    with torch.no_grad():
        approx_kl = ((ratio - 1) - logratio).mean()                               # :587
        clipfrac = ((ratio - 1.0).abs() > clip_coef).float().mean() if clip_coef is not None else torch.zeros(())
    mb_advantages = advantages
    adv_mean = adv_std = torch.zeros(())
    if norm_adv and mb_advantages.shape[0] > 1:                                   # :602-605
        adv_mean, adv_std = mb_advantages.mean(), mb_advantages.std()
        mb_advantages = (mb_advantages - adv_mean) / (adv_std + 1e-8)
    pg_loss1 = -mb_advantages * ratio                                             # :608
    if clip_coef is not None:                                                     # :609-615
        pg_loss2 = -mb_advantages * torch.clamp(ratio, 1 - clip_coef, 1 + clip_coef)
        pg_loss = torch.max(pg_loss1, pg_loss2).mean()
    else:                                                                         # :616-620
        pg_loss = pg_loss1.mean()
    espo_dev = (ratio - 1).abs().mean()
    newvalue = newvalue.view(-1, 1)                                               # :623
    target = returns.view(-1, 1)
    if critic_loss == "l2":                                                       # :632-641
        v_loss_unclipped = torch.nn.functional.mse_loss(newvalue, target)
    elif critic_loss == "l1":
        v_loss_unclipped = torch.nn.functional.l1_loss(newvalue, target)
    else:
        v_loss_unclipped = torch.nn.functional.smooth_l1_loss(newvalue, target)   # only reachable with clip_vloss
    if clip_vloss:                                                                # :665-690
        ov = old_values.view(-1, 1)
        v_clipped = ov + torch.clamp(newvalue - ov, -clip_coef, clip_coef)
        if critic_loss == "l2":
            v_loss_clipped = (v_clipped - target) ** 2
        elif critic_loss == "l1":
            v_loss_clipped = torch.abs(v_clipped - target)
        else:
            v_loss_clipped = torch.nn.functional.smooth_l1_loss(v_clipped, target, reduction="none")
        v_loss = torch.max(v_loss_unclipped, v_loss_clipped).mean()
    else:
        v_loss = v_loss_unclipped
    entropy_loss = entropy.mean()                                                 # :692
    if update <= 0:                                                               # :693-700
        target_distrib = torch.ones_like(unmasked_probs) / unmasked_probs.shape[1]
        uniform_loss = torch.nn.functional.l1_loss(unmasked_probs, target_distrib, reduction="sum")
        loss = v_loss + 0.1 * uniform_loss                                        # :701-705
    else:
        loss = pg_loss - ent_coef * entropy_loss + v_loss * vf_coef
    stats = dict(pg_loss=pg_loss, v_loss=v_loss, entropy_loss=entropy_loss, approx_kl=approx_kl, clipfrac=clipfrac,
                 espo_dev=espo_dev, adv_mean=adv_mean, adv_std=adv_std)
    return loss, stats
