import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
torch.backends.cuda.matmul.allow_tf32 = True
from wheatley_b200 import dense
M, K, HD = 5003, 64, 256
g = torch.Generator().manual_seed(9)
x = torch.randn(M, K, generator=g).cuda().requires_grad_(True)
ws = [(torch.randn(HD, K, generator=g) / 8).cuda().requires_grad_(True) for _ in range(3)]
gs = [torch.randn(M, HD, generator=g).cuda() for _ in range(3)]
outs = dense.project3(x, *ws)
torch.cuda.synchronize(); print("fwd ok")
g3 = torch.stack(gs, 0)
torch.autograd.backward(outs, (g3[0], g3[1], g3[2]))
torch.cuda.synchronize(); print("bwd ok")
