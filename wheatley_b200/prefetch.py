"""Batch prefetch for the PPO update loop: the NEXT minibatch's host->device copies, batch build, graph plan and
clique view are queued BEFORE the current minibatch's training step instead of at the start of its own step, and
the plan's one device->host read of its counts is asynchronous (``GraphPlan.prepare(defer=True)``), so by the
time the next step asks for them (``get()``) they have long arrived: the host never waits on a GPU that has only
the plan build to do (~1 ms at 8 x Taillard 100x20), and in the host-bound regime the build runs on the GPU while
the host is still launching the forward.

Two placements:

* ``side_stream=False`` (default): everything on the caller's stream, ahead of the step's kernels.  No
  cross-stream bookkeeping; right for raw observations (``submit_observation``: 2 MB of copies per minibatch).
* ``side_stream=True``: copies and kernels on a side stream, so that a LARGE copy (the 41 MB int64 COO of an
  already batched graph, ``submit``) runs on the copy engine and the batch / plan kernels run under the step's
  kernels; the two streams are ordered by one event per submit / get (``_side_waits_for_main``).  Wins when the step
  is GPU-bound (8 x Taillard 100x20: 4.6 vs 5.2 ms end to end).

The reference builds each minibatch on the host inside the update loop (alg/ppo.py:547-727 ->
jssp/models/agent_observation / gnn_dgl.py:272-286) with nothing overlapped; this is the B200-side counterpart
of a pinned-memory DataLoader with ``non_blocking`` copies.

    pf = BatchPrefetcher(device)
    pf.submit_observation(obs[0], max_n_nodes, extras[0], conflicts="clique")
    for i in range(n_minibatches):
        d, plan = pf.get()                      # ready on the current stream; d["batch"] is the JsspBatch
        if i + 1 < n_minibatches:
            pf.submit_observation(obs[i + 1], max_n_nodes, extras[i + 1], conflicts="clique")
        step(d)
"""
import contextlib

import torch

from . import _lib
from .plan import GraphPlan


class BatchPrefetcher:
    def __init__(self, device, src_key="src", dst_key="dst", etype_key="etype", side_stream=False):
        # (a worker thread for the submissions was tried: the step's own launches and the worker's contend for
        # the GIL and the step got 50 % slower; the submissions stay on the caller's thread)
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _lib.B200GatError("BatchPrefetcher needs a CUDA device: wheatley_b200 has no CPU path")
        self.side = torch.cuda.Stream(self.device) if side_stream else None
        self.keys = (src_key, dst_key, etype_key)
        self._pending = []

    def _enqueue(self, job):
        self._pending.append(job())

    def _stream(self):
        if self.side is None:
            return contextlib.nullcontext()
        self._side_waits_for_main()
        return torch.cuda.stream(self.side)

    def _side_waits_for_main(self):
        """Memory discipline of the side-stream placement.  A batch is allocated from the side stream's pool, used by
        the training step on the caller's stream and freed by reference counting once that step has been QUEUED, at
        which point the allocator may hand its blocks to the side stream's next allocation.  Everything queued on the
        side stream therefore first waits for what the caller's stream has been given so far: one event per submit /
        get instead of ``Tensor.record_stream`` on each of a batch's ~70 tensors (0.3 ms of host time per step, which
        made this placement lose whenever the loop was host-bound).  Contract: a prefetched batch is consumed on the
        stream that is current when ``get()`` is called."""
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self.side.wait_event(ev)

    def submit(self, host: dict, n_nodes: int):
        """Queue the copies of ``host`` (a dict of pinned CPU tensors) and the plan build on the side stream."""
        ks, kd, kt = self.keys

        def job():
            with self._stream():
                d = {k: v.to(self.device, non_blocking=True) for k, v in host.items()}
                plan = GraphPlan(d[ks], d[kd], n_nodes)
                plan.prepare(d.get(kt), defer=True)
                done = torch.cuda.Event()
                done.record()
            return d, plan, done
        self._enqueue(job)

    def submit_observation(self, obs: dict, max_n_nodes: int, extras: dict = None, **builder_kw):
        """Queue a raw JSSP gym observation: copies of the padded arrays, the device-side batch builder
        (``wheatley_b200.batch``), the plan and its clique view.  ``extras``: other pinned host tensors of the
        minibatch (masks, actions, advantages ...) copied alongside.  ``get()`` returns them with the built
        ``JsspBatch`` under key ``"batch"``."""
        from .batch import jssp_batch_from_observation

        def job():
            with self._stream():
                d = {k: v.to(self.device, non_blocking=True) for k, v in (extras or {}).items()}
                jb, built = jssp_batch_from_observation(obs, self.device, max_n_nodes, defer=True, **builder_kw)
                d.update(batch=jb, **{k: built[k] for k in ("src", "dst", "etype", "feat", "poolnodes")})
                done = torch.cuda.Event()
                done.record()
            return d, jb.plan, done
        self._enqueue(job)

    def submit_psp(self, agent, arena, extras: dict = None, n_graphs=None, **preprocess_kw):
        """Queue an arena of PSP wire-format observation slots (``wheatley_b200.wire``): copies, the device batch
        builder (``agent.preprocess`` = psp/models/agent.py:298-406 on the device), the plan.  ``get()`` returns the
        extras with the 9-tuple of ``preprocess`` under key ``"batch"``."""
        def job():
            with self._stream():
                d = {k: v.to(self.device, non_blocking=True) for k, v in (extras or {}).items()}
                obs = agent.preprocess(arena, n_graphs, defer=True, **preprocess_kw)
                d["batch"] = obs
                done = torch.cuda.Event()
                done.record()
            return d, obs[0].plan, done
        self._enqueue(job)

    def get(self):
        """``(device_tensors, plan)`` of the oldest submitted batch, usable on the current stream."""
        if not self._pending:
            raise _lib.B200GatError("BatchPrefetcher.get() without a submitted batch")
        d, plan, done = self._pending.pop(0)
        if self.side is None:
            plan.finish()               # waits for the counts' event only (queued ahead of the previous step)
            return d, plan
        main = torch.cuda.current_stream(self.device)
        self._side_waits_for_main()
        with torch.cuda.stream(self.side):
            plan.finish()               # allocations stay on the side pool; zero-fills the hub arrival counters
            done = torch.cuda.Event()   # (a kernel queued AFTER the event of submit time)
            done.record()
        main.wait_event(done)
        return d, plan


class HostReadback:
    """Device -> host reads of per-step scalars (loss, the PPO statistics) WITHOUT stalling the launch thread: every
    ``push`` queues an asynchronous copy into pinned memory behind the step's kernels and records an event; ``pop``
    waits for the OLDEST outstanding event only.  A loop that pushes step i and pops step i-1 reads every step's
    value, one step late, while the host is already launching the next step.  (``loss.item()`` makes the host wait
    for the whole step, so the next step's launches start on an idle GPU: at 8 x Taillard 100x20 that serialisation
    cost 1.5 ms of a 6.9 ms step.  The reference reads eight scalars per minibatch with ``.item()``,
    alg/ppo.py:588-595,707-710, for logging and the per-epoch KL early stop, neither of which needs the value before
    the next minibatch is launched.)"""

    def __init__(self, depth=2, numel=1):
        self.bufs = [torch.empty(numel, dtype=torch.float32, pin_memory=True) for _ in range(depth + 1)]
        self.events = [torch.cuda.Event() for _ in range(depth + 1)]
        self.head = self.tail = 0          # monotonically increasing push / pop counters

    def __len__(self):
        return self.head - self.tail

    def push(self, t: torch.Tensor):
        if len(self) >= len(self.bufs):
            raise _lib.B200GatError("HostReadback is full: pop() before pushing more")
        k = self.head % len(self.bufs)
        self.bufs[k][: t.numel()].copy_(t.detach().reshape(-1), non_blocking=True)
        self.events[k].record()
        self.head += 1

    def pop(self):
        if not len(self):
            raise _lib.B200GatError("HostReadback.pop() without a pending read")
        k = self.tail % len(self.bufs)
        self.events[k].synchronize()
        self.tail += 1
        return self.bufs[k].clone()
