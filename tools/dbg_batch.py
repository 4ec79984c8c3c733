"""Development probe: host and device cost of the device-side batch builder at the bench size."""
import os, sys, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import wheatley_b200 as wb
from wheatley_b200 import _lib
from wheatley_b200.batch import jssp_sizes, jssp_batch_on_device, jssp_batch_from_observation
from wheatley_b200.workload import jssp_observation_batch
obs = {k: v.contiguous().pin_memory() for k, v in jssp_observation_batch(100, 20, 8).items()}
dev = torch.device("cuda:0")


def host_t(name, fn, n=50):
    for _ in range(5): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
    print(f"{name:40s} host {1e6*(t1-t0)/n:8.1f} us/call   incl. drain {1e6*(t2-t0)/n:8.1f} us/call")


host_t("jssp_sizes (host only)", lambda: jssp_sizes(obs["features"], obs["n_nodes"], obs["n_edges"], "clique"))
host_t("jssp_batch_on_device", lambda: jssp_batch_on_device(obs, dev, "clique", "learn"))
host_t("jssp_batch_from_observation", lambda: jssp_batch_from_observation(obs, dev, 2000, "clique", "learn"))
host_t("  same, defer=True (no finish)", lambda: jssp_batch_from_observation(obs, dev, 2000, "clique", "learn", defer=True))
_lib.timing_enable(True)
for _ in range(10):
    jssp_batch_on_device(obs, dev, "clique", "learn")
torch.cuda.synchronize()
tm = _lib.timing_read()
print("builder kernels:", tm["batch"][0] / max(tm["batch"][1], 1) * 1e3, "us per call over", tm["batch"][1], "calls")
_lib.timing_enable(False)
