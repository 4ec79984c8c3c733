"""GPU parity for the tcgen05 TF32 GEMM (b200gat_gemm_tf32) and the fused dense path built on it
(GELU MLP with fused epilogues, the three-way EGAT projection), against fp64 torch references.

Tolerance: the tensor core reads fp32 operands as TF32 (10-bit mantissa, truncated), the precision the
reference runs its matmuls in (psp/train_psp.py:47).  Per GEMM: max-abs error <= 3e-3 of the max-abs
reference value (K <= 768); through the whole trunk (GEMM chains 12 deep, fwd+bwd): <= 2e-2 on the outputs
and the loss, <= 5e-2 on every parameter gradient."""
import pytest
import torch

pytestmark = pytest.mark.gpu

TOL_GEMM = 3e-3


@pytest.fixture(scope="module")
def wb():
    import wheatley_b200
    return wheatley_b200


@pytest.fixture()
def tf32():
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    yield
    torch.backends.cuda.matmul.allow_tf32 = prev


def rel(a, b):
    b = b.detach().double().cpu()
    return float((a.detach().double().cpu() - b).abs().max()) / max(float(b.abs().max()), 1e-6)


def launches():
    from wheatley_b200 import _lib
    return _lib.launch_count()


@pytest.mark.parametrize("M,K,N", [(128, 64, 64), (1000, 64, 256), (16008, 256, 256), (16008, 256, 64),
                                   (333, 464, 64), (5, 32, 128), (4100, 36, 192), (1000, 64, 100), (300, 32, 332)])
def test_linear_forward_variants(wb, tf32, M, K, N):
    from wheatley_b200 import dense
    g = torch.Generator().manual_seed(M + K + N)
    x = torch.randn(M, K, generator=g).cuda()
    W = (torch.randn(N, K, generator=g) / K ** 0.5).cuda()
    b = torch.randn(N, generator=g).cuda()
    r = torch.randn(M, N, generator=g).cuda()
    ref = x.double() @ W.double().t() + b.double()
    l0 = launches()
    _, y = dense.linear_fwd(x, W, b)
    assert launches() > l0, "the tcgen05 kernel did not run"
    assert rel(y, ref) < TOL_GEMM
    _, y = dense.linear_fwd(x, W, b, residual=r)
    assert rel(y, ref + r.double()) < TOL_GEMM
    _, y = dense.linear_fwd(x, W, None)
    assert rel(y, ref - b.double()) < TOL_GEMM
    z, h = dense.linear_fwd(x, W, b, gelu=True)
    assert rel(z, ref) < TOL_GEMM
    assert rel(h, torch.nn.functional.gelu(ref)) < TOL_GEMM
    z2, h2 = dense.linear_fwd(x, W, b, gelu=True, want_z=False)
    assert z2 is None and torch.equal(h2, h)


@pytest.mark.parametrize("M,N,K", [(128, 64, 64), (16008, 256, 256), (16008, 64, 256), (16008, 256, 64),
                                   (777, 128, 192), (16000, 64, 464), (900, 100, 36)])
def test_linear_backward_variants(wb, tf32, M, N, K):
    """dx = dy W (MN-major B), dx * gelu'(z) epilogue, dW = dy^T x (both operands MN-major, split-K),
    bias gradient column sums."""
    from wheatley_b200 import dense
    g = torch.Generator().manual_seed(M + 3 * N + K)
    dy = torch.randn(M, N, generator=g).cuda()
    W = (torch.randn(N, K, generator=g) / N ** 0.5).cuda()
    x = torch.randn(M, K, generator=g).cuda()
    z = torch.randn(M, K, generator=g).cuda()
    l0 = launches()
    dx = dense.linear_dx(dy, W)
    assert launches() > l0
    ref_dx = dy.double() @ W.double()
    assert rel(dx, ref_dx) < TOL_GEMM
    dxz = dense.linear_dx(dy, W, z)
    zd = z.double().cpu().requires_grad_(True)
    torch.nn.functional.gelu(zd).backward(ref_dx.cpu())
    assert rel(dxz, zd.grad) < TOL_GEMM
    dW, db = dense.linear_dw(dy, x, want_bias=True)
    ref_dw = dy.double().t() @ x.double()
    assert rel(dW, ref_dw) < TOL_GEMM
    assert rel(db, dy.double().sum(0)) < 1e-5              # exact fp32 row sums of the staged dy tiles
    dW2, db2 = dense.linear_dw(dy, x, want_bias=True)
    assert torch.equal(dW, dW2) and torch.equal(db, db2), "split-K must be deterministic"
    assert torch.equal(dense.linear_dw(dy, x)[0], dW)
    assert rel(dense.colsum(dy), dy.double().sum(0)) < 1e-5


def test_gemm_rejects_uncovered_shapes(wb, tf32):
    from wheatley_b200 import _lib, dense
    x = torch.randn(10, 30).cuda()           # 30-wide rows: not 16-byte aligned -> library GEMM (MLPs pad instead)
    W = torch.randn(64, 30).cuda()
    l0 = launches()
    _, y = dense.linear_fwd(x, W, None)
    assert launches() == l0
    assert rel(y, x.double() @ W.double().t()) < TOL_GEMM
    g = dense._desc(0, 0, 10, 1, 64, x, 64, W, 64, y, 1)     # N = 1
    assert _lib.lib.b200gat_gemm_supported(g) == 0
    assert _lib.lib.b200gat_gemm_tf32(g, None) != 0 and b"multiple of 4" in _lib.lib.b200gat_last_error()


@pytest.mark.parametrize("dims,res", [((256, 256, 256, 64), True), ((30, 64, 64, 64), False),
                                      ((464, 64, 64, 1), False)])
def test_gelu_mlp_matches_autograd(wb, tf32, dims, res):
    """MLP module on the fused path (mixed with library GEMMs where a layer is not covered) against
    fp64 autograd of the same module."""
    from wheatley_b200.mlp import MLP
    M = 3001
    torch.manual_seed(5)
    mlp = MLP(3, dims[0], dims[1], dims[3], False, "gelu").cuda()
    ref = MLP(3, dims[0], dims[1], dims[3], False, "gelu").double()
    ref.load_state_dict({k: v.double().cpu() for k, v in mlp.state_dict().items()})
    x = torch.randn(M, dims[0]).cuda().requires_grad_(True)
    r = torch.randn(M, dims[3]).cuda().requires_grad_(True) if res else None
    gy = torch.randn(M, dims[3]).cuda()
    l0 = launches()
    y = mlp(x, residual=r)
    y.backward(gy)
    assert launches() > l0
    xr = x.detach().double().cpu().requires_grad_(True)
    rr = r.detach().double().cpu().requires_grad_(True) if res else None
    torch.backends.cuda.matmul.allow_tf32 = False
    yr = ref._forward_library(xr)
    if res:
        yr = yr + rr
    yr.backward(gy.double().cpu())
    assert rel(y, yr) < 2 * TOL_GEMM
    assert rel(x.grad, xr.grad) < 3 * TOL_GEMM
    if res:
        assert rel(r.grad, rr.grad) < 1e-6
    for (k, p), (_, q) in zip(mlp.named_parameters(), ref.named_parameters()):
        assert rel(p.grad, q.grad) < 3 * TOL_GEMM, k


def test_project3_matches_three_linears(wb, tf32):
    from wheatley_b200 import dense
    M, K, HD = 5003, 64, 256
    g = torch.Generator().manual_seed(9)
    x = torch.randn(M, K, generator=g).cuda().requires_grad_(True)
    ws = [(torch.randn(HD, K, generator=g) / 8).cuda().requires_grad_(True) for _ in range(3)]
    gs = [torch.randn(M, HD, generator=g).cuda() for _ in range(3)]
    l0 = launches()
    outs = dense.project3(x, *ws)
    assert launches() == l0 + 1, "one launch for the three projections"
    assert all(o.is_contiguous() for o in outs)
    # gradients arriving in one [3, M, HD] buffer (what the conv backward produces) and separately
    g3 = torch.stack(gs, 0)
    for grads in ((g3[0], g3[1], g3[2]), tuple(gs)):
        x.grad = None
        for w in ws:
            w.grad = None
        torch.autograd.backward(dense.project3(x, *ws), grads)
        xd = x.detach().double()
        ref_dx = sum(gg.double() @ w.detach().double() for gg, w in zip(gs, ws))
        assert rel(x.grad, ref_dx) < TOL_GEMM
        for gg, w in zip(gs, ws):
            assert rel(w.grad, gg.double().t() @ xd) < TOL_GEMM
    for o, w in zip(outs, ws):
        assert rel(o, x.detach().double() @ w.detach().double().t()) < TOL_GEMM


def test_trunk_with_tf32_dense_path_matches_oracle(wb, tf32):
    """Whole hot path with the tcgen05 dense path on (the bench configuration): 4-graph batch through
    GnnDGL (3 layers, hidden 64) + heads + loss, fwd and bwd, against the fp64 oracle."""
    from oracle import trunk_oracle
    from wheatley_b200.agent import JsspAgent
    from wheatley_b200.gnn_dgl import JsspBatch
    from wheatley_b200.ppo import ppo_like_loss
    from wheatley_b200.workload import jssp_batch
    n_j, n_m, B, L, Hd, H = 12, 6, 4, 3, 64, 4
    hb = jssp_batch(n_j, n_m, B, seed0=3, progress=0.3)
    F = hb["feat"].shape[1]
    torch.manual_seed(1)
    agent = JsspAgent(F, n_j * n_m, n_m, Hd, L, H).cuda()
    gnn = trunk_oracle.OracleGnnDGL(F, "learn", n_j * n_m, n_m, 3, L, Hd, "gelu", H)
    ref = trunk_oracle.OracleAgent(gnn, F + Hd * (L + 1)).double()
    ref.load_state_dict({k: v.double().cpu() for k, v in agent.state_dict().items()})
    g = torch.Generator().manual_seed(1)
    masks = torch.rand(B, n_j * n_m, generator=g) < 0.4
    masks[:, 0] = True
    action = torch.multinomial(masks.float(), 1, generator=g).squeeze(1)
    adv, ret = torch.randn(B, generator=g), torch.randn(B, generator=g)
    plan = wb.GraphPlan(hb["src"].cuda(), hb["dst"].cuda(), hb["n_nodes"])
    NT = hb["n_task_nodes"]
    jb = JsspBatch(plan, hb["etype"].cuda(), hb["feat"].cuda(), B, n_j * n_m, NT,
                   poolnodes=torch.arange(NT, NT + B).cuda())
    from wheatley_b200 import _lib
    _lib.timing_enable(True)
    _, lp, ent, val, _ = agent.get_action_and_value(jb, action.cuda(), masks.cuda())
    loss = ppo_like_loss(lp, ent, val, adv.cuda(), ret.cuda())
    loss.backward()
    torch.cuda.synchronize()
    tm = _lib.timing_read()
    n_gemm = tm["gemm"][1] + tm["gemm_dx"][1] + tm["gemm_dw"][1]
    _lib.timing_enable(False)
    assert n_gemm >= L * 12, f"only {n_gemm} tcgen05 GEMM launches in the step"
    args_ = (hb["src"], hb["dst"], hb["etype"], hb["feat"].double(), B, n_j * n_m, NT)
    lp_r, ent_r, val_r = ref.evaluate(args_, action, masks)
    loss_r = trunk_oracle.ppo_like_loss(lp_r, ent_r, val_r, adv.double(), ret.double())
    loss_r.backward()
    assert rel(lp, lp_r) < 2e-2 and rel(ent, ent_r) < 2e-2 and rel(val, val_r) < 2e-2
    assert rel(loss, loss_r) < 2e-2
    refp = dict(ref.named_parameters())
    for k, p in agent.named_parameters():
        if refp[k].grad is None:
            continue
        assert rel(p.grad, refp[k].grad) < 5e-2, k


def test_graphconv_pyg_flavour_with_tf32_projections(wb, tf32):
    """GraphConv(pyg=True) -> GATv2Conv: lin_l / lin_r run on the tcgen05 GEMM when TF32 is allowed; outputs
    and gradients against the fp64 oracle (psp/graph/graph_conv.py:33-40,51-53)."""
    from oracle import gat_oracle
    from wheatley_b200 import _lib
    N, E, Fin, D, H = 3000, 20000, 64, 64, 4
    gen = torch.Generator().manual_seed(11)
    src = torch.randint(0, N, (E,), generator=gen)
    dst = torch.randint(0, N, (E,), generator=gen)
    ei = torch.stack([torch.cat([src, torch.arange(N)]), torch.cat([dst, torch.arange(N)])])
    ref = gat_oracle.OracleGATv2Conv(Fin, D, H, edge_dim=Fin).double()
    with torch.no_grad():
        for p_ in ref.parameters():
            p_.add_(0.05 * torch.randn(p_.shape, generator=gen, dtype=torch.float64))
    layer = wb.GraphConv(Fin, D, H, pyg=True, edge_dim=Fin).cuda()
    layer.conv.load_state_dict({k: v.float() for k, v in ref.state_dict().items()})
    x = torch.randn(N, Fin, generator=gen, dtype=torch.float64)
    ea = torch.randn(ei.shape[1], Fin, generator=gen, dtype=torch.float64)
    R = torch.randn(N, H * D, generator=gen, dtype=torch.float64)
    xr = x.clone().requires_grad_(True)
    out_ref, _ = gat_oracle.gatv2_forward(xr, ei, ea, ref.lin_l.weight, ref.lin_l.bias, ref.lin_r.weight,
                                          ref.lin_r.bias, ref.lin_edge.weight, ref.att, ref.bias)
    (out_ref * R).sum().backward()
    xg = x.float().cuda().requires_grad_(True)
    _lib.timing_enable(True)
    out, _ = layer.forward_nog(xg, ei.cuda(), ea.float().cuda())
    (out * R.float().cuda()).sum().backward()
    torch.cuda.synchronize()
    tm = _lib.timing_read()
    _lib.timing_enable(False)
    assert tm["gemm"][1] >= 2 and tm["gemm_dx"][1] >= 2 and tm["gemm_dw"][1] >= 2, tm
    assert rel(out, out_ref) < 1e-2
    assert rel(xg.grad, xr.grad) < 5e-2          # TF32 logits through the softmax: same bound as the trunk test
    for k, p_ in layer.conv.named_parameters():
        q = dict(ref.named_parameters())[k]
        assert rel(p_.grad, q.grad) < 5e-2, k


@pytest.mark.parametrize("M,K,N", [(128, 64, 64), (16008, 256, 256), (16008, 64, 768), (333, 464, 64), (4100, 36, 192)])
def test_fp32x3_mode_has_fp32_matmul_accuracy(wb, M, K, N):
    """``allow_tf32 = False``: b200gat_gemm_tf32 runs 3xTF32 split arithmetic (B200GAT_GEMM_FP32X3, three
    tcgen05.mma per k step on hi / lo operand tiles).  Forward (K-major operands), dx (MN-major B) and dW (both
    MN-major, split-K, bias row sums) against fp64: <= 1e-5 of the largest entry (measured 3.5e-6; an fp32 matmul
    gives 1-2e-6, the TF32 path 1-3e-3)."""
    from wheatley_b200 import _lib, dense
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        g = torch.Generator().manual_seed(M + K + N)
        x = torch.randn(M, K, generator=g).cuda()
        W = (torch.randn(N, K, generator=g) / K ** 0.5).cuda()
        b = torch.randn(N, generator=g).cuda()
        dy = torch.randn(M, N, generator=g).cuda()
        l0 = launches()
        _, y = dense.linear_fwd(x, W, b)
        assert launches() > l0 and _lib.lib.b200gat_gemm_precision_get() == _lib.GEMM_FP32X3
        assert rel(y, x.double() @ W.double().t() + b.double()) < 1e-5
        z, h = dense.linear_fwd(x, W, b, gelu=True)
        assert rel(h, torch.nn.functional.gelu(x.double() @ W.double().t() + b.double())) < 1e-5
        dx = dense.linear_dx(dy, W)
        assert rel(dx, dy.double() @ W.double()) < 1e-5
        dW, db = dense.linear_dw(dy, x, want_bias=True)
        assert rel(dW, dy.double().t() @ x.double()) < 1e-5
        assert rel(db, dy.double().sum(0)) < 1e-5
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
