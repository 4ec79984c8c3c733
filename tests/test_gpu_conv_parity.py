"""GPU parity: the CUDA graph-attention layers (through the C-ABI library) against the CPU
oracle in float64 and against the golden vectors produced by executing the reference.

Tolerance (north_star): fp32 kernels within 1e-4 relative error of the fp64 oracle, measured
as max|got-ref| / max|ref| per tensor (``REL``); integer outputs bit-exact."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

REL = 1e-4


def rel_err(got, ref):
    got = got.detach().double().cpu()
    ref = ref.detach().double().cpu()
    assert got.shape == ref.shape, (got.shape, ref.shape)
    denom = max(float(ref.abs().max()), 1e-30)
    return float((got - ref).abs().max()) / denom


def assert_close(got, ref, what, rel=REL, atol=2e-6):
    """max|got-ref| <= rel * max|ref| + atol (atol covers gradients that are identically zero in
    exact arithmetic, e.g. the softmax-invariant attention bias of GATv3)."""
    got = got.detach().double().cpu()
    ref = ref.detach().double().cpu()
    assert got.shape == ref.shape, (what, got.shape, ref.shape)
    err = float((got - ref).abs().max()) if ref.numel() else 0.0
    bound = rel * (float(ref.abs().max()) if ref.numel() else 0.0) + atol
    assert err <= bound, f"{what}: abs err {err:.3e} > {bound:.3e} (rel {rel:.0e})"


@pytest.fixture(scope="module")
def wb():
    import wheatley_b200
    return wheatley_b200


def random_graph(rng, n, e, self_loops=True, hub=None):
    src = rng.integers(0, n, e)
    dst = rng.integers(0, n, e)
    if hub is not None:                   # one very high in-degree / out-degree node
        k = e // 3
        dst[:k] = hub
        src[k:2 * k] = hub
    if self_loops:
        src = np.concatenate([src, np.arange(n)])
        dst = np.concatenate([dst, np.arange(n)])
    perm = rng.permutation(src.shape[0])
    return torch.from_numpy(src[perm]).long(), torch.from_numpy(dst[perm]).long()


def banded_graph(rng, n, self_loops=True):
    """The structure the windowed kernels (csrc/gat_win.cu) are selected for (E <= 8 N): job chains of 10 nodes with
    forward and reversed precedence edges, self loops, and a pool node (the last one: a hub when n > 257) wired to
    and from every other node."""
    i = np.arange(n - 2)
    i = i[(i + 1) % 10 != 0]
    pool = n - 1
    others = np.arange(n - 1)
    src = np.concatenate([i, i + 1, others, np.full(n - 1, pool)])
    dst = np.concatenate([i + 1, i, np.full(n - 1, pool), others])
    if self_loops:
        src = np.concatenate([src, np.arange(n)])
        dst = np.concatenate([dst, np.arange(n)])
    perm = rng.permutation(src.shape[0])
    return torch.from_numpy(src[perm]).long(), torch.from_numpy(dst[perm]).long()


def _graph_arg(wb, src, dst, N, graph, edge_index=None):
    """What the layer is called with: the COO as the reference passes it, or - banded graphs - a prebuilt plan
    flagged B200GAT_GRAPH_LOW_DEGREE, which selects the windowed kernels (the plan sets the flag itself only when no
    adjacency list is longer than 32; here the pool node's is, on purpose: hub CTAs, multi-chunk segments)."""
    if graph != "banded":
        return edge_index.cuda() if edge_index is not None else (src.cuda(), dst.cuda(), N)
    from wheatley_b200 import _lib
    plan = wb.GraphPlan(src.cuda(), dst.cuda(), N)
    plan.c.flags |= _lib.GRAPH_LOW_DEGREE
    return plan


def load_state(dst_mod, src_mod):
    with torch.no_grad():
        sd = {k: v.detach().float() for k, v in src_mod.state_dict().items()}
        dst_mod.load_state_dict(sd)


def perturb(mod, gen):
    with torch.no_grad():
        for p in mod.parameters():
            p.add_(0.1 * torch.randn(p.shape, generator=gen, dtype=p.dtype))


def grid(t, q):
    """Round to multiples of 1/q in [-2, 2].  With node / edge features on a 1/8 grid and the weights that feed the
    pre-activation on a 1/32 grid, every pre-activation z = Pi[u] + Pj[v] + Pe (+ b) is computed EXACTLY in fp32 by
    any correct implementation (all partial sums fit 24 bits), so kernel and fp64 oracle take the same side of the
    leaky_relu kink on every element.  With unrestricted random inputs about one pre-activation in 1e6 lands within
    fp32 rounding of 0 and takes the other slope depending on how the projections were rounded (library fp32 GEMM,
    3xTF32 tcgen05 GEMM and the oracle differ at the 1e-7 level), which moves single gradient entries by up to 1e-3
    of the maximum - a property of evaluating the layer in fp32, not of a kernel (same device as
    tests/test_gpu_clique.py)."""
    return torch.round(t.clamp(-2, 2) * q) / q


def grid_params(mod, names):
    with torch.no_grad():
        for k, p in mod.named_parameters():
            if k in names:
                p.copy_(grid(p, 32))


GRAD_X_REL = REL

CASES = [  # N, E, Fin, D, H
    (17, 60, 6, 5, 3),
    (300, 4000, 64, 64, 4),
    (64, 3000, 32, 32, 4),
    (50, 400, 16, 128, 2),
    (40, 300, 8, 16, 1),
]


BANDED_CASES = [  # N, Fin, D, H: H*D = 256 / 128 / 64; 403 and 150 are not multiples of the 16-node window
    (403, 16, 64, 4),
    (403, 16, 32, 4),
    (150, 8, 16, 4),
    (1040, 8, 64, 4),
]


@pytest.mark.parametrize("N,E,Fin,D,H", CASES)
@pytest.mark.parametrize("efmode", ["dense", "typed"])
def test_egat_matches_oracle(wb, N, E, Fin, D, H, efmode):
    _egat_vs_oracle(wb, N, E, Fin, D, H, efmode, "random")


@pytest.mark.parametrize("N,Fin,D,H", BANDED_CASES)
def test_egat_banded_graph_matches_oracle(wb, N, Fin, D, H):
    """Low-degree banded graph with a hub: the windowed kernels (k_fwd_win / k_bwd_dst_win / k_bwd_src_win) + the hub
    CTAs of the warp-per-node kernels."""
    _egat_vs_oracle(wb, N, 0, Fin, D, H, "typed", "banded")


def _egat_vs_oracle(wb, N, E, Fin, D, H, efmode, graph):
    from oracle import gat_oracle
    rng = np.random.default_rng(N * 7 + E)
    gen = torch.Generator().manual_seed(N + E)
    src, dst = random_graph(rng, N, E, hub=3) if graph == "random" else banded_graph(rng, N)
    Et = src.shape[0]
    ref = gat_oracle.OracleEGATConv(Fin, Fin, D, D, H).double()
    perturb(ref, gen)
    grid_params(ref, ("fc_ni.weight", "fc_nj.weight", "fc_fij.weight", "bias"))
    layer = wb.EGATConv(Fin, Fin, D, D, H).cuda()
    load_state(layer, ref)
    x = grid(torch.randn(N, Fin, generator=gen, dtype=torch.float64), 8)
    T = 7
    etype = torch.from_numpy(rng.integers(0, T, Et)).long()
    table = grid(torch.randn(T, Fin, generator=gen, dtype=torch.float64), 8)
    ef_dense = table[etype] if efmode == "typed" else grid(torch.randn(Et, Fin, generator=gen, dtype=torch.float64), 8)
    R = torch.randn(N, H, D, generator=gen, dtype=torch.float64)

    xr = x.clone().requires_grad_(True)
    efr = (table if efmode == "typed" else ef_dense).clone().requires_grad_(True)
    h_ref, f_ref, a_ref = ref(src, dst, xr, efr[etype] if efmode == "typed" else efr, get_attention=True)
    (h_ref * R).sum().backward()

    xg = x.float().cuda().requires_grad_(True)
    efg = (table if efmode == "typed" else ef_dense).float().cuda().requires_grad_(True)
    efeats = wb.TypedEdgeFeats(etype.cuda(), efg) if efmode == "typed" else efg
    h, f, a = layer(_graph_arg(wb, src, dst, N, graph), xg, efeats, get_attention=True)
    (h * R.float().cuda()).sum().backward()

    assert_close(h, h_ref, "h_out")
    assert_close(a, a_ref, "attention")
    assert_close(f.materialize(), f_ref, "f_out")
    assert_close(xg.grad, xr.grad, "grad x", rel=GRAD_X_REL)
    assert_close(efg.grad, efr.grad, "grad efeats")
    for (k, p), (_, pr) in zip(layer.named_parameters(), ref.named_parameters()):
        assert_close(p.grad, pr.grad, "grad " + k)


def test_egat_golden(wb, golden_dir):
    g = np.load(os.path.join(golden_dir, "eegat.npz"))
    T = lambda k: torch.from_numpy(g[k])
    H = int(g["heads"])
    Fin, Fe, D = g["x"].shape[1], g["ef"].shape[1], g["p.attn"].shape[2]
    layer = wb.EGATConv(Fin, Fe, D, D, H).cuda()
    with torch.no_grad():
        layer.load_state_dict({k[2:]: T(k).float() for k in g.files if k.startswith("p.")})
    x = T("x").float().cuda().requires_grad_(True)
    ef = T("ef").float().cuda().requires_grad_(True)
    h, f, a = layer((T("src").cuda(), T("dst").cuda(), x.shape[0]), x, ef, get_attention=True)
    (h * T("R").float().cuda()).sum().backward()
    assert_close(h, T("h"), "h")
    assert_close(f.materialize(), T("f"), "f")
    assert_close(a, T("a"), "a")
    assert_close(x.grad, T("gx"), "gx")
    assert_close(ef.grad, T("gef"), "gef")
    for k, p in layer.named_parameters():
        assert_close(p.grad, T("g." + k), "g." + k)


def test_eegat_edge_scoring_golden(wb, golden_dir):
    g = np.load(os.path.join(golden_dir, "eegat_edge_scoring.npz"))
    T = lambda k: torch.from_numpy(g[k])
    H = int(g["heads"])
    Fin, Fe, D = g["x"].shape[1], g["ef"].shape[1], g["p.attn"].shape[2]
    layer = wb.EEGATConv(Fin, Fe, D, D, H, edge_scoring=True).cuda()
    with torch.no_grad():
        layer.load_state_dict({k[2:]: T(k).float() for k in g.files if k.startswith("p.")})
    h, f, e_out, a = layer((T("src").cuda(), T("dst").cuda(), g["x"].shape[0]), T("x").float().cuda(),
                           T("ef").float().cuda(), T("efe").float().cuda(), get_attention=True)
    assert_close(h, T("h"), "h")
    assert_close(f.materialize(), T("f"), "f")
    assert_close(e_out, T("e_out"), "e_out")
    assert_close(a, T("a"), "a")


def test_egat_zero_in_degree_raises(wb):
    layer = wb.EGATConv(4, 4, 3, 3, 2).cuda()
    x = torch.randn(3, 4).cuda()
    with pytest.raises(wb.DGLError):
        layer((torch.tensor([0, 1]).cuda(), torch.tensor([1, 1]).cuda(), 3), x, torch.randn(2, 4).cuda())


@pytest.mark.parametrize("N,E,Fin,D,H", CASES)
@pytest.mark.parametrize("efmode", ["dense", "affine", "typed", "none"])
def test_gatv2_matches_oracle(wb, N, E, Fin, D, H, efmode):
    _gatv2_vs_oracle(wb, N, E, Fin, D, H, efmode, "random")


@pytest.mark.parametrize("N,Fin,D,H", BANDED_CASES)
@pytest.mark.parametrize("efmode", ["affine", "none"])
def test_gatv2_banded_graph_matches_oracle(wb, N, Fin, D, H, efmode):
    """V aliasing Pi and the affine edge term on the windowed kernels."""
    _gatv2_vs_oracle(wb, N, 0, Fin, D, H, efmode, "banded")


def _gatv2_vs_oracle(wb, N, E, Fin, D, H, efmode, graph):
    from oracle import gat_oracle
    rng = np.random.default_rng(N * 3 + E)
    gen = torch.Generator().manual_seed(N * 5 + E)
    if graph == "random":
        src, dst = random_graph(rng, N, E, self_loops=False, hub=1)    # isolated destinations allowed
    else:
        src, dst = banded_graph(rng, N, self_loops=False)
    ei = torch.stack([src, dst])
    Et = src.shape[0]
    ref = gat_oracle.OracleGATv2Conv(Fin, D, H, edge_dim=Fin).double()
    perturb(ref, gen)
    grid_params(ref, ("lin_l.weight", "lin_l.bias", "lin_r.weight", "lin_r.bias", "lin_edge.weight"))
    layer = wb.GATv2Conv(Fin, D, H, add_self_loops=False, edge_dim=Fin).cuda()
    load_state(layer, ref)
    x = grid(torch.randn(N, Fin, generator=gen, dtype=torch.float64), 8)
    R = torch.randn(N, H * D, generator=gen, dtype=torch.float64)
    Tn, K = 16, 5
    etype = torch.from_numpy(rng.integers(0, Tn, Et)).long()
    table = grid(torch.randn(Tn, Fin, generator=gen, dtype=torch.float64), 8)
    attrs = grid(torch.randn(Et, K, generator=gen, dtype=torch.float64), 8)
    attrs[torch.rand(Et, generator=gen) < 0.5] = 0.0
    aw = grid(torch.randn(Fin, K, generator=gen, dtype=torch.float64), 8)
    leaves_ref = [t.clone().requires_grad_(True) for t in (table, aw)]
    leaves_gpu = [t.float().cuda().requires_grad_(True) for t in (table, aw)]
    if efmode == "dense":
        ea_ref = grid(torch.randn(Et, Fin, generator=gen, dtype=torch.float64), 8).requires_grad_(True)
        ea_gpu = ea_ref.detach().float().cuda().requires_grad_(True)
        leaves_ref, leaves_gpu = [ea_ref], [ea_gpu]
    elif efmode == "affine":
        ea_ref = leaves_ref[0][etype] + attrs @ leaves_ref[1].t()
        ea_gpu = wb.AffineEdgeFeats(etype.cuda(), leaves_gpu[0], attrs.float().cuda(), leaves_gpu[1])
    elif efmode == "typed":
        ea_ref = leaves_ref[0][etype]
        ea_gpu = wb.TypedEdgeFeats(etype.cuda(), leaves_gpu[0])
        leaves_ref, leaves_gpu = leaves_ref[:1], leaves_gpu[:1]
    else:
        ea_ref = ea_gpu = None
        leaves_ref, leaves_gpu = [], []
    xr = x.clone().requires_grad_(True)
    out_ref, alpha_ref = gat_oracle.gatv2_forward(
        xr, ei, ea_ref, ref.lin_l.weight, ref.lin_l.bias, ref.lin_r.weight, ref.lin_r.bias,
        ref.lin_edge.weight, ref.att, ref.bias)
    (out_ref * R).sum().backward()
    xg = x.float().cuda().requires_grad_(True)
    out, (_, alpha) = layer(xg, _graph_arg(wb, src, dst, N, graph, ei), ea_gpu, return_attention_weights=True)
    (out * R.float().cuda()).sum().backward()
    assert_close(out, out_ref, "out")
    assert_close(alpha, alpha_ref, "alpha")
    assert_close(xg.grad, xr.grad, "grad x", rel=GRAD_X_REL)
    for i, (a, b) in enumerate(zip(leaves_gpu, leaves_ref)):
        assert_close(a.grad, b.grad, f"grad edge leaf {i}")
    for (k, p), (_, pr) in zip(layer.named_parameters(), ref.named_parameters()):
        if efmode == "none" and k.startswith("lin_edge"):
            assert p.grad is None or not p.grad.abs().any()
            continue
        assert_close(p.grad, pr.grad, "grad " + k)


def test_gatv2_isolated_destination_gets_bias(wb):
    layer = wb.GATv2Conv(4, 3, 2, add_self_loops=False, edge_dim=4).cuda()
    with torch.no_grad():
        layer.bias.copy_(torch.arange(6.0))
    x = torch.randn(3, 4).cuda()
    out = layer(x, torch.tensor([[0, 1], [1, 1]]).cuda(), torch.randn(2, 4).cuda())
    assert torch.equal(out[2].cpu(), torch.arange(6.0))
    assert torch.equal(out[0].cpu(), torch.arange(6.0))


def test_graphconv_boundary(wb):
    """psp/graph/graph_conv.py:51-60: (out [N, H*D], None) from forward and forward_nog."""
    from oracle import gat_oracle

    class G:
        pass
    torch.manual_seed(0)
    N, E, F, D, H = 30, 200, 16, 16, 4
    gc = wb.GraphConv(F, D, H, pyg=True).cuda()
    assert isinstance(gc.conv, wb.GATv2Conv) and gc.gatv3 is False
    assert sorted(k for k, _ in gc.named_parameters()) == sorted(
        "conv." + k for k in ("att", "lin_l.weight", "lin_l.bias", "lin_r.weight", "lin_r.bias",
                              "lin_edge.weight", "bias"))
    g = G()
    g.edge_index = torch.randint(0, N, (2, E)).cuda()
    x, ea = torch.randn(N, F).cuda(), torch.randn(E, F).cuda()
    out, none = gc(g, x, ea)
    out2, none2 = gc.forward_nog(x, g.edge_index, ea)
    assert none is None and none2 is None and out.shape == (N, H * D)
    assert torch.equal(out, out2)
    ref = gat_oracle.OracleGATv2Conv(F, D, H, edge_dim=F).double()
    ref.load_state_dict({k[5:]: v.double().cpu() for k, v in gc.state_dict().items()})
    assert_close(out, ref(x.double().cpu(), g.edge_index.cpu(), ea.double().cpu()), "GraphConv out")
    out3, _ = gc.forward_nog(x, g.edge_index, None)       # gnn_hier.py:69,115
    assert_close(out3, ref(x.double().cpu(), g.edge_index.cpu(), None), "edge_attr=None")
    # DGL flavour through the same adapter: (f.flatten(-2,-1), ef)
    gd = wb.GraphConv(F, D, H, pyg=False).cuda()
    ei = torch.cat([g.edge_index, torch.arange(N).cuda().repeat(2, 1)], dim=1)
    ea2 = torch.randn(ei.shape[1], F).cuda()
    f, ef = gd((ei[0], ei[1], N), x, ea2)
    assert f.shape == (N, H * D) and tuple(ef.shape) == (ei.shape[1], H, D)


def test_gatv3_golden(wb, golden_dir):
    g = np.load(os.path.join(golden_dir, "gatv3.npz"))
    T = lambda k: torch.from_numpy(g[k])
    H = int(g["heads"])
    Fin = g["x"].shape[1]
    C = g["out"].shape[1] // H
    layer = wb.GATv3Conv(Fin, C, H, dropout=0, edge_dim=Fin, bias=True).cuda()
    assert "bias" not in layer.state_dict()            # reference quirk
    with torch.no_grad():
        layer.load_state_dict({k[2:]: T(k).float() for k in g.files if k.startswith("p.")})
    x = T("x").float().cuda().requires_grad_(True)
    out, (ei, alpha) = layer(x, T("edge_index").cuda(), T("ea").float().cuda(), return_attention_weights=True)
    (out * T("R").float().cuda()).sum().backward()
    assert torch.equal(ei.cpu(), T("edge_index_out"))
    assert_close(out, T("out"), "out")
    assert_close(alpha, T("alpha"), "alpha")
    assert_close(x.grad, T("gx"), "gx")
    for k, p in layer.named_parameters():
        if k == "lin_edge.weight":
            assert p.grad is None
            continue
        assert_close(p.grad, T("g." + k), "g." + k)


@pytest.mark.parametrize("N,E,Fin,D,H", CASES[:3])
def test_gatv3_matches_oracle(wb, N, E, Fin, D, H):
    from oracle import gat_oracle
    rng = np.random.default_rng(N + 11 * E)
    gen = torch.Generator().manual_seed(N + E + 1)
    src, dst = random_graph(rng, N, E, self_loops=False)
    ei = torch.stack([src, dst])
    ref = gat_oracle.OracleGATv3Conv(Fin, D, H, edge_dim=Fin).double()
    perturb(ref, gen)
    layer = wb.GATv3Conv(Fin, D, H, edge_dim=Fin).cuda()
    load_state(layer, ref)
    x = torch.randn(N, Fin, generator=gen, dtype=torch.float64)
    R = torch.randn(N, H * D, generator=gen, dtype=torch.float64)
    xr = x.clone().requires_grad_(True)
    (ref(xr, ei) * R).sum().backward()
    xg = x.float().cuda().requires_grad_(True)
    out = layer(xg, ei.cuda())
    (out * R.float().cuda()).sum().backward()
    assert_close(out, ref(x, ei), "out")
    assert_close(xg.grad, xr.grad, "grad x", rel=GRAD_X_REL)
    for (k, p), (_, pr) in zip(layer.named_parameters(), ref.named_parameters()):
        if k == "lin_edge.weight":
            continue
        assert_close(p.grad, pr.grad, "grad " + k)


def test_empty_graph(wb):
    layer = wb.GATv2Conv(4, 3, 2, add_self_loops=False, edge_dim=4).cuda()
    x = torch.randn(5, 4).cuda()
    out = layer(x, torch.zeros((2, 0), dtype=torch.long).cuda(), torch.zeros((0, 4)).cuda())
    assert out.shape == (5, 6) and torch.equal(out.cpu(), layer.bias.detach().cpu().expand(5, 6))
    out0 = layer(torch.zeros((0, 4)).cuda(), torch.zeros((2, 0), dtype=torch.long).cuda(), None)
    assert out0.shape == (0, 6)


def test_cpu_tensors_fail_loudly(wb):
    layer = wb.GATv2Conv(4, 3, 2, add_self_loops=False, edge_dim=4)
    with pytest.raises(wb.B200GatError):
        layer(torch.randn(5, 4), torch.zeros((2, 3), dtype=torch.long), None)


def test_full_size_properties(wb):
    """Taillard 100x20-sized clique graph (2,001 nodes, ~206k edges): size-independent
    properties - attention rows sum to one per destination, aggregation is linear in the value
    tensor, result is invariant under a permutation of the edge list, run-to-run bit-reproducible."""
    from wheatley_b200.workload import jssp_clique_graph
    src, dst, etype, n = jssp_clique_graph(100, 20, seed=0)
    src, dst, etype = src.cuda(), dst.cuda(), etype.cuda()
    E = src.shape[0]
    assert n == 2001 and E == 205801
    torch.manual_seed(0)
    H, D = 4, 64
    layer = wb.EGATConv(D, D, D, D, H).cuda()
    x = torch.randn(n, D).cuda()
    table = torch.randn(6, D).cuda()
    with torch.no_grad():
        h, _, a = layer((src, dst, n), x, wb.TypedEdgeFeats(etype, table), get_attention=True)
        sums = torch.zeros(n, H, device="cuda", dtype=torch.float64).index_add_(0, dst, a.squeeze(-1).double())
        assert float((sums - 1).abs().max()) < 1e-5
        h_again, _ = layer((src, dst, n), x, wb.TypedEdgeFeats(etype, table))
        assert torch.equal(h, h_again)
        perm = torch.randperm(E, device="cuda")
        h_perm, _ = layer((src[perm], dst[perm], n), x, wb.TypedEdgeFeats(etype[perm], table))
        assert_close(h_perm, h, "edge permutation invariance", rel=1e-5)
        # linearity in V with the logits held fixed: scale fc_node_src by 3
        layer.fc_node_src.weight.mul_(3.0)
        h3, _ = layer((src, dst, n), x, wb.TypedEdgeFeats(etype, table))
        assert_close(h3, 3.0 * h, "linearity in V", rel=1e-5)
