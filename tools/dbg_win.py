"""Windowed vs warp-per-node sparse kernels on the fused-trunk test case: determinism run to run, fused vs layer path,
and (second invocation with B200GAT_WIN=0) against the other kernel family.  Usage: python tools/dbg_win.py TAG"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
tag = sys.argv[1]
tf32 = len(sys.argv) > 2 and sys.argv[2] == "tf32"
torch.backends.cuda.matmul.allow_tf32 = tf32
from wheatley_b200.batch import jssp_batch_from_observation  # noqa: E402
from wheatley_b200.gnn_dgl import GnnDGL  # noqa: E402
from wheatley_b200.workload import jssp_observation_batch  # noqa: E402

n_j, n_m, B, L, Hd, H = 12, 5, 3, 3, 64, 4
obs = jssp_observation_batch(n_j, n_m, B, seed0=4, progress=0.4)
Fdim = obs["features"].shape[2]
torch.manual_seed(2)
gnn = GnnDGL(Fdim, "gatv2", "learn", False, n_j * n_m, n_m, 3, L, Hd, "gelu", H, residual=True, normalize=False,
             conflicts="node", mid_in_edges=True, split_output=True).cuda()
gen = torch.Generator().manual_seed(0)
R = None


def run(mode):
    global R
    os.environ["WHEATLEY_B200_TRUNK"] = mode
    for p in gnn.parameters():
        p.grad = None
    jb, _ = jssp_batch_from_observation(obs, "cuda", n_j * n_m, conflicts="node", graph_pooling="learn",
                                        mid_in_edges=True, max_n_machines=n_m)
    nf, ge = gnn(jb)
    if R is None:
        R = (torch.randn(nf.shape, generator=gen).cuda(), torch.randn(ge.shape, generator=gen).cuda())
    ((nf * R[0]).sum() + (ge * R[1]).sum()).backward()
    torch.cuda.synchronize()
    return {"nf": nf.detach().cpu(), **{k: p.grad.detach().cpu() for k, p in gnn.named_parameters() if p.grad is not None}}


def rel(a, b):
    return float((a.double() - b.double()).abs().max()) / max(float(b.double().abs().max()), 1e-6)


a1, a2, f1 = run("layers"), run("layers"), run("fused")
worst_rr = max((rel(a1[k], a2[k]), k) for k in a1)
worst_fl = max((rel(f1[k], a1[k]), k) for k in a1)
print(tag, "tf32" if tf32 else "fp32x3", "run-to-run", worst_rr, "| fused vs layers", worst_fl)
torch.save(a1, f"gpurun_out/dbg_win_{tag}.pt")
other = f"gpurun_out/dbg_win_{'win0' if tag != 'win0' else 'win1'}.pt"
if os.path.exists(other):
    b = torch.load(other)
    print(tag, "vs other kernel family", max((rel(a1[k], b[k]), k) for k in a1))
