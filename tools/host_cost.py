"""Development probe: host-side cost per call (launch only, no sync) of the library entry points vs torch ops."""
import os, sys, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
torch.backends.cuda.matmul.allow_tf32 = True
from wheatley_b200 import dense, _lib
from wheatley_b200.mlp import MLP
M = 2048
x = torch.randn(M, 256).cuda(); W = torch.randn(256, 256).cuda(); b = torch.randn(256).cuda()
y = torch.empty(M, 256).cuda()


def t(name, fn, n=2000):
    for _ in range(50): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    dt = (time.perf_counter() - t0) / n * 1e6
    torch.cuda.synchronize()
    print(f"{name:34s} {dt:7.2f} us/call")


g = dense._desc(0, 0, M, 256, 256, x, 256, W, 256, y, 256, bias=b)
sp = _lib.stream_ptr()
t("raw b200gat_gemm_tf32 (ctypes)", lambda: _lib.lib.b200gat_gemm_tf32(g, sp))
t("dense._desc", lambda: dense._desc(0, 0, M, 256, 256, x, 256, W, 256, y, 256, bias=b))
t("dense.linear_fwd", lambda: dense.linear_fwd(x, W, b))
t("torch F.linear", lambda: torch.nn.functional.linear(x, W, b))
t("torch.empty", lambda: torch.empty((M, 256), dtype=torch.float32, device=x.device))
t("_lib.stream_ptr", _lib.stream_ptr)
mlp = MLP(3, 256, 256, 64, False, "gelu").cuda()
with torch.no_grad():
    t("MLP forward no_grad (3 GEMMs)", lambda: mlp(x), 500)
xg = x.clone().requires_grad_(True)
t("MLP forward autograd", lambda: mlp(xg), 500)
def fb():
    mlp(xg).sum().backward()
t("MLP fwd+bwd", fb, 300)
