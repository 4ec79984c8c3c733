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

    def reduce(self):
        if not (dist.is_available() and dist.is_initialized()):
            return self.flat
        world = dist.get_world_size(self.group)
        if world <= 1:
            return self.flat
        self.flat = self.flat_grads()
        dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=self.group)
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
