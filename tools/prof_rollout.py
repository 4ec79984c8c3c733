"""Where a rollout (inference) step of 8 x Taillard 100x20 spends its time: observation -> device batch + plan (one
host sync for the plan's counts) -> trunk + heads under no_grad -> sampled action read back."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from wheatley_b200 import _lib  # noqa: E402
from wheatley_b200.agent import JsspAgent  # noqa: E402
from wheatley_b200.batch import jssp_batch_from_observation  # noqa: E402
from wheatley_b200.workload import jssp_observation_batch  # noqa: E402

torch.backends.cuda.matmul.allow_tf32 = True
dev = torch.device("cuda", 0)
n_j, n_m, B = 100, 20, 8
torch.manual_seed(0)
agent = JsspAgent(6 + n_m + 4, n_j * n_m, n_m, 64, 6, 4).to(dev)
obs = {k: v.contiguous().pin_memory() for k, v in jssp_observation_batch(n_j, n_m, B).items()}
masks = (torch.rand(B, n_j * n_m) < 0.3)
masks[:, 0] = True
masks = masks.pin_memory()
T = {"build": 0.0, "forward_launch": 0.0, "forward_wait": 0.0, "total": 0.0}
n = 0
for it in range(60):
    t0 = time.perf_counter()
    m = masks.to(dev, non_blocking=True)
    jb, _ = jssp_batch_from_observation(obs, dev, n_j * n_m, conflicts="clique", graph_pooling="learn")
    t1 = time.perf_counter()
    l0 = _lib.launch_count()
    with torch.no_grad():
        action, logprob, _, value, _ = agent.get_action_and_value(jb, None, m)
    t2 = time.perf_counter()
    a = action.cpu()
    t3 = time.perf_counter()
    if it >= 20:
        T["build"] += t1 - t0; T["forward_launch"] += t2 - t1; T["forward_wait"] += t3 - t2; T["total"] += t3 - t0
        n += 1
print({k: round(v / n * 1e3, 3) for k, v in T.items()}, "ms per rollout step;", _lib.launch_count() - l0, "library launches in the forward")
# device time of the forward alone
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
jb, _ = jssp_batch_from_observation(obs, dev, n_j * n_m, conflicts="clique", graph_pooling="learn")
m = masks.to(dev)
torch.cuda.synchronize()
e0.record()
with torch.no_grad():
    for _ in range(10):
        agent.get_action_and_value(jb, None, m)
e1.record()
torch.cuda.synchronize()
print("forward alone, back to back:", round(e0.elapsed_time(e1) / 10, 3), "ms")
e0.record()
for _ in range(10):
    jssp_batch_from_observation(obs, dev, n_j * n_m, conflicts="clique", graph_pooling="learn")
e1.record()
torch.cuda.synchronize()
print("batch + plan build alone (with its host sync):", round(e0.elapsed_time(e1) / 10, 3), "ms")
