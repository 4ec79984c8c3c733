"""The slice of alg/ppo.py the hot path runs under: per-minibatch loss evaluation
(alg/ppo.py:575-705) and, new in this implementation, data-parallel sharding with one NCCL
all-reduce of the flat gradient per optimizer step (the reference has no
distributed code).

Graphs of a batch are independent (block-diagonal adjacency), so ranks take contiguous shards
of whole graphs and exchange nothing on the data path."""
import torch
import torch.distributed as dist


def ppo_like_loss(logprob, entropy, value, advantages, returns, ent_coef=0.05, vf_coef=2.0):
    """``-(adv * ratio).mean() - ent_coef * entropy.mean() + vf_coef * 0.5 * mse(value, returns)``
    with ``ratio = exp(logprob - old_logprob)``, ``old_logprob = logprob.detach()`` (first epoch,
    alg/ppo.py:582-600; clipping is inactive at ratio 1)."""
    ratio = torch.exp(logprob - logprob.detach())
    pg = -(advantages * ratio).mean()
    v = 0.5 * ((value.reshape(-1) - returns) ** 2).mean()
    return pg - ent_coef * entropy.mean() + vf_coef * v


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

    def __init__(self, params, group=None):
        self.params = [p for p in params if p.requires_grad]
        self.group = group
        self.flat = None

    def zero(self):
        for p in self.params:
            p.grad = None

    def reduce(self):
        if not (dist.is_available() and dist.is_initialized()):
            return self.flat
        world = dist.get_world_size(self.group)
        if world <= 1:
            return self.flat
        live = [p for p in self.params if p.grad is not None]
        self.flat = torch.cat([p.grad.reshape(-1) for p in live])
        dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=self.group)
        self.flat.div_(world)
        off = 0
        for p in live:
            n = p.numel()
            p.grad = self.flat[off: off + n].view_as(p)
            off += n
        return self.flat
