"""Development probe: error of the MN-major GEMM variants under descriptor-field overrides."""
import os, subprocess, sys, json
CASES = {
    "default": {},
    "sw128_l2": {"B200GAT_MN_TMA_SWIZZLE": "3", "B200GAT_MN_LAYOUT": "2", "B200GAT_MN_SBO": "1024"},
    "swap": {"B200GAT_MN_LBO": "512", "B200GAT_MN_SBO": "4096"},
    "sbo1024": {"B200GAT_MN_SBO": "1024"},
    "flip8": {"B200GAT_MN_TMA_SWIZZLE": "5"},
}
if len(sys.argv) > 1 and sys.argv[1] == "child":
    import torch
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    torch.backends.cuda.matmul.allow_tf32 = True
    from wheatley_b200 import dense
    g = torch.Generator().manual_seed(0)
    out = {}
    for (M, N, K) in [(128, 64, 64), (1000, 256, 128)]:
        dy = torch.randn(M, N, generator=g).cuda(); W = torch.randn(N, K, generator=g).cuda(); x = torch.randn(M, K, generator=g).cuda()
        def rel(a, b): return float((a.double() - b).abs().max() / b.abs().max())
        out[f"dx{M}"] = rel(dense.linear_dx(dy, W), dy.double() @ W.double())
        out[f"dw{M}"] = rel(dense.linear_dw(dy, x)[0], dy.double().t() @ x.double())
    print(json.dumps(out))
else:
    for name, env in CASES.items():
        e = dict(os.environ); e.update(env)
        r = subprocess.run([sys.executable, __file__, "child"], env=e, capture_output=True, text=True, timeout=300)
        print(name, r.stdout.strip()[-300:], r.stderr.strip()[-300:] if r.returncode else "")
