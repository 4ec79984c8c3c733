"""Which objects of one PSP / JSSP batch build are only reclaimable by the cyclic GC (should be none)."""
import collections
import gc
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
dev = torch.device("cuda", 0)
from wheatley_b200.psp_batch import preprocess  # noqa: E402
from wheatley_b200.wire import PspArena  # noqa: E402
from wheatley_b200.workload import psp_taillard_graph  # noqa: E402

arena = PspArena.from_graphs([psp_taillard_graph(4, 4, b, progress=0.0) for b in range(64)])
for conflicts in ("node", "clique"):
    obs = preprocess(arena, dev, conflicts=conflicts, graph_pooling="learn")
    del obs
    gc.collect()
    gc.set_debug(gc.DEBUG_SAVEALL)
    obs = preprocess(arena, dev, conflicts=conflicts, graph_pooling="learn", defer=True)
    obs[0].plan.finish()
    del obs
    n = gc.collect()
    kinds = collections.Counter(type(o).__name__ for o in gc.garbage)
    print(conflicts, "unreachable objects found by the cyclic GC:", n, dict(kinds.most_common(8)))
    for o in gc.garbage:
        if type(o).__name__ in ("GraphPlan", "CliqueView", "PspGraphData", "dict") and not isinstance(o, dict):
            print("   ", type(o).__name__, [type(r).__name__ for r in gc.get_referrers(o)][:6])
    gc.set_debug(0)
    gc.garbage.clear()
