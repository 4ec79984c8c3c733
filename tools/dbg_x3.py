"""Development probe: accuracy of the 3xTF32 GEMM mode on the projection shapes of the attention layers."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from wheatley_b200 import dense, _lib

torch.backends.cuda.matmul.allow_tf32 = False
rel = lambda a, b: float((a.double().cpu() - b.cpu()).abs().max()) / max(float(b.abs().max()), 1e-9)
for M in (300, 333, 128, 16008, 44):
    for trial in range(3):
        g = torch.Generator().manual_seed(M + trial)
        x = torch.randn(M, 64, generator=g).cuda().requires_grad_(True)
        ws = [(torch.randn(256, 64, generator=g) / 8).cuda().requires_grad_(True) for _ in range(3)]
        R = [torch.randn(M, 256, generator=g).cuda() for _ in range(3)]
        outs = dense.project3(x, *ws)
        sum((o * r).sum() for o, r in zip(outs, R)).backward()
        xd = x.detach().double().cpu()
        e_f = max(rel(o.detach(), xd @ w.detach().double().cpu().t()) for o, w in zip(outs, ws))
        e_dx = rel(x.grad, sum(r.double().cpu() @ w.detach().double().cpu() for r, w in zip(R, ws)))
        e_dw = max(rel(w.grad, r.double().cpu().t() @ xd) for r, w in zip(R, ws))
        # single linear with bias (GATv2 lin_l)
        x2 = x.detach().clone().requires_grad_(True)
        w2, b2 = ws[0].detach().clone().requires_grad_(True), torch.randn(256, generator=g).cuda().requires_grad_(True)
        y2 = dense.dense_linear(x2, w2, b2)
        (y2 * R[0]).sum().backward()
        e2 = (rel(y2.detach(), xd @ w2.detach().double().cpu().t() + b2.detach().double().cpu()),
              rel(x2.grad, R[0].double().cpu() @ w2.detach().double().cpu()),
              rel(w2.grad, R[0].double().cpu().t() @ xd), rel(b2.grad, R[0].double().cpu().sum(0)))
        print(f"M={M} trial {trial}: proj3 fwd {e_f:.2e} dx {e_dx:.2e} dW {e_dw:.2e} | linear fwd {e2[0]:.2e} dx {e2[1]:.2e} "
              f"dW {e2[2]:.2e} db {e2[3]:.2e}  mode {_lib.lib.b200gat_gemm_precision_get()}")
