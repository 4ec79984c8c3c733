import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
import wheatley_b200 as wb
from wheatley_b200.workload import jssp_batch
from test_gpu_clique import _egat_case
nj, nm, b = [int(x) for x in (sys.argv[1:4] if len(sys.argv) > 3 else (10, 10, 3))]
hb = jssp_batch(nj, nm, b, seed0=11, progress=0.5)
def rel(a, b):
    a = a.detach().double().cpu(); b = b.detach().double().cpu()
    return float((a - b).abs().max()) / max(float(b.abs().max()), 1e-30)
for seed in range(6):
    torch.manual_seed(seed)
    gen = torch.Generator().manual_seed(nj * 100 + nm)
    res, plan = _egat_case(wb, hb, 4, 64, gen)
    errs = {k: rel(*res[k]) for k in ("h", "a", "gx", "gt")}
    errs.update({k: rel(g, gr) for k, g, gr in res["params"]})
    worst = max(errs, key=errs.get)
    print(seed, "worst", worst, "%.2e" % errs[worst], " ".join("%s=%.1e" % kv for kv in errs.items()))
