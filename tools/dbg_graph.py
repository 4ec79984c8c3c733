import sys, os, traceback
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import wheatley_b200 as wb
from wheatley_b200 import _lib
from wheatley_b200.agent import JsspAgent
from wheatley_b200.gnn_dgl import JsspBatch
from wheatley_b200.ppo import FlatGradAllReduce, ppo_like_loss
from wheatley_b200.workload import jssp_batch
dev = torch.device("cuda", 0)
nj, nm, B = 15, 15, 4
hb = jssp_batch(nj, nm, B)
torch.manual_seed(0)
agent = JsspAgent(hb["feat"].shape[1], nj * nm, nm, 64, 2, 4).to(dev)
bucket = FlatGradAllReduce(list(agent.parameters()))
opt = torch.optim.Adam(agent.parameters(), lr=2e-4, fused=True, capturable=True)
d = {k: hb[k].to(dev) for k in ("src", "dst", "etype", "feat")}
plan = wb.GraphPlan(d["src"], d["dst"], hb["n_nodes"])
NT = hb["n_task_nodes"]
jb = JsspBatch(plan, d["etype"], d["feat"], B, nj * nm, NT, poolnodes=torch.arange(NT, NT + B, device=dev))
masks = torch.ones(B, nj * nm, dtype=torch.bool, device=dev)
action = torch.zeros(B, dtype=torch.long, device=dev)
adv = torch.randn(B, device=dev); ret = torch.randn(B, device=dev)
def fwd():
    return agent.get_action_and_value(jb, action, masks)
def step(with_opt=True):
    bucket.zero()
    _, lp, ent, val, _ = fwd()
    loss = ppo_like_loss(lp, ent, val, adv, ret)
    loss.backward()
    if with_opt:
        opt.step()
    return loss
for _ in range(3):
    step()
torch.cuda.synchronize()
mode = sys.argv[1] if len(sys.argv) > 1 else "thread_local"
for name, fn in (("fwd", lambda: fwd()), ("fwd+bwd", lambda: step(False)), ("full", lambda: step(True))):
    try:
        g = torch.cuda.CUDAGraph()
        with torch.no_grad() if name == "fwd" else torch.enable_grad():
            with torch.cuda.graph(g, capture_error_mode=mode):
                fn()
        g.replay(); torch.cuda.synchronize()
        print(name, "captured OK")
    except Exception:
        print(name, "FAILED")
        traceback.print_exc(limit=12)
        torch.cuda.synchronize()
        break
