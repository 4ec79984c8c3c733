"""Device memory of the e2e loop step by step (allocated / reserved MiB, cudaMalloc calls): a plan <-> view reference
cycle used to leave every minibatch's plan to the cyclic GC.  Usage: python tools/dbg_leak.py [psp272|taillard100x20]"""
import gc
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
wl = sys.argv[1] if len(sys.argv) > 1 else "psp272"
dev = torch.device("cuda", 0)
import wheatley_b200 as wb  # noqa: E402
from wheatley_b200 import _lib  # noqa: E402

gc.collect()
gc.freeze()
gc.set_debug(0)
n_gc = [0]
gc.callbacks.append(lambda phase, info: n_gc.__setitem__(0, n_gc[0] + (phase == "start" and info["generation"] == 2)))
if wl == "psp272":
    from wheatley_b200.psp_batch import preprocess
    from wheatley_b200.wire import PspArena
    from wheatley_b200.workload import psp_taillard_graph
    arena = PspArena.from_graphs([psp_taillard_graph(4, 4, b, progress=0.0) for b in range(1536)])
    build = lambda: preprocess(arena, dev, conflicts="node", graph_pooling="learn", defer=True)[0].plan
else:
    from wheatley_b200.batch import jssp_batch_from_observation
    from wheatley_b200.workload import jssp_observation_batch
    obs = {k: v.contiguous().pin_memory() for k, v in jssp_observation_batch(100, 20, 8, seed0=0, progress=0.0).items()}
    build = lambda: jssp_batch_from_observation(obs, dev, 2000, conflicts="clique", graph_pooling="learn", defer=True)[0].plan
plans = []
for step in range(61):
    plans.append(build())
    if len(plans) > 2:
        plans.pop(0).finish()
    if step % 10 == 0:
        torch.cuda.synchronize()
        st = torch.cuda.memory_stats(dev)
        print(wl, "step", step, "allocated MiB", torch.cuda.memory_allocated(dev) >> 20, "reserved MiB",
              torch.cuda.memory_reserved(dev) >> 20, "cudaMalloc", st["num_device_alloc"], "gen2 collections", n_gc[0])
