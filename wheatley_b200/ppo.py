"""The slice of alg/ppo.py the hot path runs under: per-minibatch loss evaluation
(alg/ppo.py:575-705) and, new in this implementation, data-parallel sharding with one NCCL
all-reduce of the flat gradient per optimizer step (the reference has no
distributed code).

Graphs of a batch are independent (block-diagonal adjacency), so ranks take contiguous shards
of whole graphs and exchange nothing on the data path."""
import ctypes as C

import torch
import torch.distributed as dist

from . import _lib
from ._lib import eager_only


def ppo_like_loss(logprob, entropy, value, advantages, returns, ent_coef=0.05, vf_coef=2.0):
    """``-(adv * ratio).mean() - ent_coef * entropy.mean() + vf_coef * 0.5 * mse(value, returns)``
    with ``ratio = exp(logprob - old_logprob)``, ``old_logprob = logprob.detach()`` (first epoch,
    alg/ppo.py:582-600; clipping is inactive at ratio 1)."""
    ratio = torch.exp(logprob - logprob.detach())
    pg = -(advantages * ratio).mean()
    v = 0.5 * ((value.reshape(-1) - returns) ** 2).mean()
    return pg - ent_coef * entropy.mean() + vf_coef * v


@eager_only
def gae(rewards, values, dones, next_value, next_done, gamma, gae_lambda, sigma=1.0):
    """``(advantages, returns)`` of alg/ppo.py:276-299 on the device (``b200gat_gae``): ``rewards`` / ``values``
    ``[T, E, R]``, ``dones [T, E]``, ``next_value [E, R]``, ``next_done [E]``; bit-exact against the reference's fp32
    loop over the steps."""
    if not rewards.is_cuda:
        raise _lib.B200GatError("wheatley_b200 kernels need CUDA tensors (no CPU fallback)")
    T, E, R = rewards.shape
    f = lambda t: t.to(torch.float32).contiguous()
    rewards, values, dones, next_value, next_done = f(rewards), f(values), f(dones), f(next_value), f(next_done)
    adv, ret = torch.empty_like(rewards), torch.empty_like(rewards)
    with _lib.device_guard(rewards.device):
        _lib.check(_lib.lib.b200gat_gae(_lib.ptr(rewards), _lib.ptr(values), _lib.ptr(dones), _lib.ptr(next_value),
                                        _lib.ptr(next_done), T, E, R, gamma, gae_lambda, sigma, _lib.ptr(adv),
                                        _lib.ptr(ret), _lib.stream_ptr()), "b200gat_gae")
    return adv, ret


STATS = ("pg_loss", "v_loss", "entropy_loss", "approx_kl", "clipfrac", "espo_dev", "adv_mean", "adv_std")


class _PpoLoss(torch.autograd.Function):
    """Scalar loss + statistics from one launch; the gradients of the three differentiable inputs are produced by
    the same launch, so ``backward`` is three scalings."""

    @staticmethod
    def forward(ctx, logprob, entropy, value, old_logprob, advantages, returns, old_values, moments, cfg):
        if not logprob.is_cuda:
            raise _lib.B200GatError("wheatley_b200 kernels need CUDA tensors (no CPU fallback)")
        f = lambda t: None if t is None else t.detach().to(torch.float32).reshape(-1).contiguous()
        lp, ent, val = f(logprob), f(entropy), f(value)
        olp, adv, ret, ov = f(old_logprob), f(advantages), f(returns), f(old_values)
        n = lp.numel()
        for t in (ent, val, olp, adv, ret, ov):
            if t is not None and t.numel() != n:
                raise _lib.B200GatError("PPO loss inputs must have one entry per sample")
        dev = lp.device
        out = torch.empty(1 + len(STATS), dtype=torch.float32, device=dev)
        g = torch.empty((3, n), dtype=torch.float32, device=dev)
        a = _lib.PpoLossArgs(n, _lib.ptr(lp), _lib.ptr(olp), _lib.ptr(adv), _lib.ptr(ent), _lib.ptr(val),
                             _lib.ptr(ret), _lib.ptr(ov), -1.0 if cfg["clip_coef"] is None else cfg["clip_coef"],
                             int(cfg["norm_adv"]), int(cfg["clip_vloss"]), _lib.CRITIC[cfg["critic_loss"]],
                             cfg["ent_coef"], cfg["vf_coef"], _lib.ptr(moments))
        with _lib.device_guard(dev):
            _lib.check(_lib.lib.b200gat_ppo_loss(C.byref(a), out.data_ptr(), out.data_ptr() + 4, g[0].data_ptr(),
                                                 None if ent is None else g[1].data_ptr(), g[2].data_ptr(),
                                                 _lib.stream_ptr()), "b200gat_ppo_loss")
        ctx.save_for_backward(g)
        ctx.shapes = (logprob.shape, None if entropy is None else entropy.shape, value.shape)
        stats = out[1:]
        ctx.mark_non_differentiable(stats)
        return out[0], stats

    @staticmethod
    def backward(ctx, g_loss, _g_stats):
        (g,) = ctx.saved_tensors
        s_lp, s_ent, s_val = ctx.shapes
        need = ctx.needs_input_grad
        if need[0] and need[2] and (s_ent is None or need[1]):
            g = g * g_loss                                # one launch for all three
            return (g[0].view(s_lp), None if s_ent is None else g[1].view(s_ent), g[2].view(s_val),
                    None, None, None, None, None, None)
        return ((g[0] * g_loss).view(s_lp) if need[0] else None,
                (g[1] * g_loss).view(s_ent) if (need[1] and s_ent is not None) else None,
                (g[2] * g_loss).view(s_val) if need[2] else None, None, None, None, None, None, None)


@eager_only
def ppo_loss(logprob, entropy, value, old_logprob, advantages, returns, old_values=None, *, clip_coef=0.2,
             norm_adv=True, clip_vloss=False, critic_loss="l2", ent_coef=0.0, vf_coef=0.5, update=1,
             unmasked_probs=None, group=None, sharded=False):
    """The minibatch loss of alg/ppo.py:582-705: ``(loss, stats)`` with ``stats`` a device tensor of ``STATS`` -
    what the reference reads back with eight ``.item()`` calls per minibatch stays on the device until the caller
    wants it (one copy per epoch for the KL early stop, :741-749).

    ``advantages``: already aggregated over reward components (``agent.aggregate_reward``, :599-601).
    ``update == 0``: the reference's warm-up objective ``v_loss + 0.1 * l1(unmasked probs, uniform)`` (:693-705),
    needs ``unmasked_probs [B, n]``.
    ``sharded=True`` (the reference has no distributed code): the minibatch is split over the ranks of
    ``group``; the advantage mean / std come from one all-reduce of three doubles, loss / stats / gradients are this
    rank's share of the GLOBAL mean, so the gradient exchange is a plain SUM and the result is the one a single
    process computes on the whole minibatch, however unequal the shards."""
    cfg = dict(clip_coef=clip_coef, norm_adv=norm_adv, clip_vloss=clip_vloss, critic_loss=critic_loss,
               ent_coef=ent_coef if update > 0 else 0.0, vf_coef=vf_coef)
    moments = None
    if sharded and dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        moments = adv_moments(advantages)
        dist.all_reduce(moments, op=dist.ReduceOp.SUM, group=group)
    if update <= 0:                                                        # :693-705 warm-up objective
        if unmasked_probs is None:
            raise ValueError("the warm-up objective (update == 0) needs the unmasked action probabilities")
        # v_loss + 0.1 * uniform_loss: the policy terms are evaluated for the statistics only (no gradient)
        loss, stats = _PpoLoss.apply(logprob.detach(), None if entropy is None else entropy.detach(), value, old_logprob, advantages, returns, old_values,
                                     moments, cfg)
        uniform = (unmasked_probs - 1.0 / unmasked_probs.shape[1]).abs().sum()
        return (loss - stats[0]) / vf_coef + 0.1 * uniform, stats
    loss, stats = _PpoLoss.apply(logprob, entropy, value, old_logprob, advantages, returns, old_values, moments, cfg)
    return loss, stats


@eager_only
def adv_moments(advantages):
    """``double[3] {count, sum, sum of squares}`` of this rank's advantages (``b200gat_ppo_adv_moments``)."""
    a = advantages.detach().to(torch.float32).reshape(-1).contiguous()
    if not a.is_cuda:                                                      # gloo tests of the host logic
        a64 = a.double()
        return torch.stack((torch.tensor(float(a.numel()), dtype=torch.float64), a64.sum(), (a64 * a64).sum()))
    m = torch.empty(3, dtype=torch.float64, device=a.device)
    with _lib.device_guard(a.device):
        _lib.check(_lib.lib.b200gat_ppo_adv_moments(_lib.ptr(a), a.numel(), _lib.ptr(m), _lib.stream_ptr()),
                   "b200gat_ppo_adv_moments")
    return m


def shard_range(n_items: int, rank: int, world: int):
    """Contiguous shard [lo, hi) of ``n_items`` graphs / envs for ``rank`` (sizes differ by <= 1)."""
    base, rem = divmod(n_items, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_by_edges(edges_per_graph, world: int):
    """Split whole graphs into ``world`` contiguous shards balanced by edge count (SURVEY 8e:
    "balance by sum E_g, not by graph count").  Returns a list of (lo, hi)."""
    total = float(sum(edges_per_graph))
    bounds, acc, lo = [], 0.0, 0
    n = len(edges_per_graph)
    for r in range(world):
        target = total * (r + 1) / world
        hi = lo
        while hi < n and (acc + edges_per_graph[hi] <= target + 1e-9 or hi == lo) and (n - hi) > (world - 1 - r):
            acc += edges_per_graph[hi]
            hi += 1
        if r == world - 1:
            hi = n
        bounds.append((lo, hi))
        lo = hi
    return bounds


class FlatGradAllReduce:
    """Data-parallel gradient exchange: ONE ``all_reduce(SUM)`` of the flat fp32 gradient per optimizer
    step (NCCL over NVLink on GPUs, gloo in the CPU tests).

    ``zero()`` drops the ``.grad`` tensors instead of zero-filling them, so the backward pass *moves* every
    parameter gradient into place (no per-parameter accumulate kernel: ~100 launches per step on this
    model).  ``reduce()`` concatenates the gradients into one flat buffer (one kernel), all-reduces it,
    scales by 1 / world and re-points every ``.grad`` at its slice of the buffer.  With one rank there is
    nothing to exchange and the gradients stay where the backward put them."""

    ALIGN = 64      # elements: every parameter's slot starts on a 256-byte boundary (TMA operands need 16)

    def __init__(self, params, group=None):
        self.params = [p for p in params if p.requires_grad]
        self.group = group
        self.flat = None
        self._zeros = {}
        self.offsets, off = [], 0
        for p in self.params:
            self.offsets.append(off)
            off += (p.numel() + self.ALIGN - 1) // self.ALIGN * self.ALIGN
        self.numel = off

    def _zero(self, n, like):
        z = self._zeros.get((n, like.device, like.dtype))
        if z is None:
            z = self._zeros[(n, like.device, like.dtype)] = torch.zeros(n, dtype=like.dtype, device=like.device)
        return z

    def zero(self):
        for p in self.params:
            p.grad = None

    def flat_grads(self):
        """All parameters' gradients as ONE flat tensor in parameter order, each slot padded to ``ALIGN`` elements
        (``offsets``; one ``cat`` kernel); a parameter without a gradient contributes zeros."""
        parts = []
        for p in self.params:
            n = p.numel()
            parts.append(p.grad.reshape(-1) if p.grad is not None else self._zero(n, p))
            pad = -n % self.ALIGN
            if pad:
                parts.append(self._zero(pad, p))
        return torch.cat(parts)

    def reduce(self, average=True, n_local=None):
        """``average=True``: mean of the ranks' gradients (every rank evaluated the mean loss of an EQUAL shard).
        ``n_local``: this rank's sample count - the ranks' gradients are then weighted by their share of the global
        minibatch (the gradient of the global-batch mean for unequal shards, ``shard_by_edges``): the count rides
        in the padding behind the last parameter, so it costs no extra collective.
        ``average=False``: plain SUM, for losses that are already this rank's share of the global mean
        (``ppo_loss(..., sharded=True)``)."""
        if not (dist.is_available() and dist.is_initialized()):
            return self.flat
        world = dist.get_world_size(self.group)
        if world <= 1:
            return self.flat
        self.flat = self.flat_grads()
        if n_local is not None:
            self.flat = torch.cat((self.flat * float(n_local), self.flat.new_tensor([float(n_local)])))
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=self.group)
            self.flat = self.flat[:-1] / self.flat[-1]
        else:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=self.group)
            if average:
                self.flat.div_(world)
        for p, off in zip(self.params, self.offsets):
            p.grad = self.flat[off: off + p.numel()].view_as(p)
        return self.flat


class FlatAdam:
    """Adam over ONE flat parameter (the optimizer step of alg/ppo.py:724-727 on the all-reduced
    bucket).  The parameters are re-pointed at slices of one flat buffer, the gradients are concatenated (one
    kernel; shared with ``FlatGradAllReduce`` when the ranks exchange them anyway) and ``torch.optim.Adam``
    (fused) runs on the single flat tensor: one launch over 1.7 M elements instead of multi-tensor launches over
    ~100 small ones (3 x 30 us of GPU time and ~250 us of host time per step on this model).  Element for element
    the arithmetic is ``torch.optim.Adam``'s.  A parameter without a gradient contributes zeros - the same as
    skipping it, as long as a parameter that ever gets a gradient gets one every step (true for a fixed model)."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, bucket=None):
        self.params = [p for p in params if p.requires_grad]
        self.bucket = bucket if bucket is not None else FlatGradAllReduce(self.params)
        if [id(p) for p in self.bucket.params] != [id(p) for p in self.params]:
            raise ValueError("FlatAdam and its gradient bucket must cover the same parameters in the same order")
        ref = self.params[0]
        flat = torch.zeros(self.bucket.numel, dtype=ref.dtype, device=ref.device)
        for p, off in zip(self.params, self.bucket.offsets):
            flat[off: off + p.numel()].copy_(p.detach().reshape(-1))
        self.flat = torch.nn.Parameter(flat)
        for p, off in zip(self.params, self.bucket.offsets):          # aligned slots (kernels take 16-byte pointers)
            p.data = self.flat.data[off: off + p.numel()].view_as(p)
        on_gpu = flat.is_cuda
        self.inner = torch.optim.Adam([self.flat], lr=lr, betas=betas, eps=eps, weight_decay=weight_decay,
                                      fused=on_gpu, capturable=on_gpu)

    def zero_grad(self, set_to_none=True):
        self.bucket.zero()

    def step(self):
        """After ``backward()`` (and ``bucket.reduce()`` on several ranks)."""
        world = dist.get_world_size(self.bucket.group) if dist.is_available() and dist.is_initialized() else 1
        g = self.bucket.flat if (world > 1 and self.bucket.flat is not None) else self.bucket.flat_grads()
        self.flat.grad = g
        self.inner.step()
        self.flat.grad = None
        self.bucket.flat = None

    def state_dict(self):
        return self.inner.state_dict()

    def load_state_dict(self, sd):
        self.inner.load_state_dict(sd)

    def reference_state_dict(self):
        """Adam state in ``torch.optim.Adam``'s per-parameter layout (what the reference's ``optimizer.pkl`` holds and
        ``optimizer.load_state_dict`` restores, alg/ppo.py:447-461): the flat moments sliced at the bucket offsets."""
        st = self.inner.state.get(self.flat, {})
        out = {"state": {}, "param_groups": [dict(self.inner.param_groups[0], params=list(range(len(self.params))))]}
        if st:
            for i, (p, off) in enumerate(zip(self.params, self.bucket.offsets)):
                sl = slice(off, off + p.numel())
                out["state"][i] = {"step": st["step"].detach().clone().cpu(),
                                   "exp_avg": st["exp_avg"][sl].view_as(p).clone(),
                                   "exp_avg_sq": st["exp_avg_sq"][sl].view_as(p).clone()}
        return out

    def load_reference_state_dict(self, sd):
        """Inverse of ``reference_state_dict``: resume from a reference optimizer checkpoint."""
        if not sd.get("state"):
            return
        if self.flat not in self.inner.state or not self.inner.state[self.flat]:
            self.flat.grad = torch.zeros_like(self.flat)          # materialise the state tensors
            lr = self.inner.param_groups[0]["lr"]
            self.inner.param_groups[0]["lr"] = 0.0
            self.inner.step()
            self.inner.param_groups[0]["lr"] = lr
            self.flat.grad = None
        st = self.inner.state[self.flat]
        st["exp_avg"].zero_()
        st["exp_avg_sq"].zero_()
        for i, (p, off) in enumerate(zip(self.params, self.bucket.offsets)):
            e = sd["state"].get(i)
            if e is None:
                continue
            sl = slice(off, off + p.numel())
            st["exp_avg"][sl].copy_(e["exp_avg"].reshape(-1))
            st["exp_avg_sq"][sl].copy_(e["exp_avg_sq"].reshape(-1))
            st["step"].fill_(float(e["step"]))
