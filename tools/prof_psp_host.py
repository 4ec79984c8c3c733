"""cProfile of the host side of the psp272 end-to-end step (1,536 wire-format observations per minibatch; the loop is
host-bound: 5.7 ms per step for 2.8 ms of device work).  Prints the 30 most expensive functions by cumulative time."""
import cProfile
import os
import pstats
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from wheatley_b200 import ppo  # noqa: E402
from wheatley_b200.agent import PspAgent  # noqa: E402
from wheatley_b200.prefetch import BatchPrefetcher, HostReadback  # noqa: E402
from wheatley_b200.wire import PspArena  # noqa: E402
from wheatley_b200.workload import psp_taillard_graph  # noqa: E402

torch.backends.cuda.matmul.allow_tf32 = True
dev = torch.device("cuda", 0)
B, n_j, n_m = 1536, 4, 4
max_n = n_j * n_m + 2
arena = PspArena.from_graphs([psp_taillard_graph(n_j, n_m, b, progress=0.0) for b in range(B)])
torch.manual_seed(0)
agent = PspAgent(max_n, n_m, hidden_dim=32, n_layers=4, conflicts="node").to(dev)
bucket = ppo.FlatGradAllReduce(list(agent.parameters()))
opt = ppo.FlatAdam(agent.parameters(), lr=2e-4, bucket=bucket)
g = torch.Generator().manual_seed(1)
masks = torch.rand(B, max_n, generator=g) < 0.3
masks[:, 0] = True
ex = dict(masks=masks, action=torch.multinomial(masks.float(), 1, generator=g).squeeze(1), adv=torch.randn(B, generator=g),
          ret=torch.randn(B, generator=g), old_lp=-torch.rand(B, generator=g) - 4, old_v=torch.randn(B, generator=g))
ex = {k: v.contiguous().pin_memory() for k, v in ex.items()}
kw = dict(clip_coef=0.2, norm_adv=True, critic_loss="l2", ent_coef=0.05, vf_coef=0.5)
pf = BatchPrefetcher(dev)
rb = HostReadback(depth=3)
pf.submit_psp(agent, arena, ex)
pf.submit_psp(agent, arena, ex)


def step():
    d, _ = pf.get()
    pf.submit_psp(agent, arena, ex)
    bucket.zero()
    loss, _ = agent.ppo_minibatch(d["batch"], d["action"], d["masks"], d["old_lp"], d["adv"], d["ret"], d["old_v"], **kw)
    loss.backward()
    bucket.reduce()
    opt.step()
    rb.push(loss)
    if len(rb) > 2:
        rb.pop()


for _ in range(20):
    step()
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(50):
    step()
torch.cuda.synchronize()
pr.disable()
st = pstats.Stats(pr)
st.sort_stats("cumulative").print_stats(32)
