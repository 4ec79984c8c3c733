"""Development probe: which aten ops (with shapes) run in one eager training step, forward and backward -
to find zero-fills / copies that the module code causes."""
import collections, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torch.profiler import profile, ProfilerActivity
import wheatley_b200 as wb
from wheatley_b200.agent import JsspAgent
from wheatley_b200.batch import jssp_batch_from_observation
from wheatley_b200.ppo import FlatGradAllReduce, ppo_like_loss
from wheatley_b200.workload import jssp_observation_batch
torch.backends.cuda.matmul.allow_tf32 = True
dev = torch.device("cuda", 0)
n_j, n_m, B = 100, 20, 8
agent = JsspAgent(6 + n_m + 4, n_j * n_m, n_m, 64, 6, 4).to(dev)
bucket = FlatGradAllReduce(list(agent.parameters()))
opt = torch.optim.Adam(agent.parameters(), lr=2e-4, fused=True, capturable=True)
obs = {k: v.pin_memory() for k, v in jssp_observation_batch(n_j, n_m, B).items()}
g = torch.Generator().manual_seed(1)
masks = (torch.rand(B, n_j * n_m, generator=g) < 0.3); masks[:, 0] = True
action = torch.multinomial(masks.float(), 1, generator=g).squeeze(1).to(dev)
masks = masks.to(dev); adv = torch.randn(B, generator=g).to(dev); ret = torch.randn(B, generator=g).to(dev)


def step():
    jb, _ = jssp_batch_from_observation(obs, dev, n_j * n_m, conflicts="clique", graph_pooling="learn")
    bucket.zero()
    _, lp, ent, val, _ = agent.get_action_and_value(jb, action, masks)
    loss = ppo_like_loss(lp, ent, val, adv, ret)
    loss.backward()
    bucket.reduce()
    opt.step()
    return loss.item()


for _ in range(3):
    step()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], record_shapes=True, with_stack=True) as prof:
    step()
torch.cuda.synchronize()
cnt = collections.Counter()
for e in prof.events():
    if e.name in ("aten::fill_", "aten::zero_", "aten::copy_", "aten::clone", "aten::contiguous", "aten::cat",
                  "aten::add", "aten::add_", "aten::constant_pad_nd", "aten::zeros", "aten::zeros_like",
                  "aten::index_put_", "aten::sum", "aten::mm", "aten::addmm"):
        st = [f for f in (e.stack or []) if "wheatley_b200" in f or "bench" in f or "prof_ops" in f]
        cnt[(e.name, str(e.input_shapes)[:60], (st[0].split("/")[-1][:60] if st else "autograd/none"))] += 1
for (name, shp, where), c in sorted(cnt.items(), key=lambda x: (x[0][0], -x[1])):
    print(f"{c:4d} {name:22s} {shp:62s} {where}")
