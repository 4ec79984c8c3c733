"""GPU tests of the dense clique-block path: exact block detection / residual CSR (integer work,
bit-exact against a numpy restatement) and parity of the clique kernels with the fp64 oracle on
schedule graphs (forward, attention weights, every gradient), EGAT and GATv2 flavours.

Tolerance: 1e-4 relative (max|got-ref| / max|ref|), the same bound as the sparse kernels; the
tensor-core aggregation uses 3xTF32 split operands to stay inside it."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

REL = 1e-4


def assert_close(got, ref, what, rel=REL, atol=2e-6):
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


def np_blocks(src, dst, etype, n_nodes, min_size=8, max_size=128):
    """numpy restatement of b200gat_clique_build's definition: per node the longest run (first on ties)
    of consecutive destination-sorted in-edges with one type and strictly ascending sources != v; nodes
    whose (run + self) sets coincide and whose runs list exactly the other members form a block."""
    order = np.lexsort((np.arange(src.shape[0]), dst))      # stable by dst
    s_sorted, t_sorted = src[order], etype[order]
    ptr = np.concatenate([[0], np.cumsum(np.bincount(dst, minlength=n_nodes))])
    runs = {}
    for v in range(n_nodes):
        b, e = ptr[v], ptr[v + 1]
        best = (0, b)
        cur = b
        for p in range(b, e + 1):
            brk = p == e or p == b or t_sorted[p] != t_sorted[p - 1] or s_sorted[p] <= s_sorted[p - 1] \
                or s_sorted[p] == v or s_sorted[p - 1] == v
            if brk:
                if p - cur > best[0]:
                    best = (p - cur, cur)
                cur = p + 1 if (p < e and s_sorted[p] == v) else p
        ln, rb = best
        if min_size - 1 <= ln <= max_size - 1:
            runs[v] = (rb, ln, int(t_sorted[rb]), tuple(s_sorted[rb:rb + ln]))
    groups = {}
    for v, (rb, ln, t, srcs) in runs.items():
        key = (tuple(sorted(srcs + (v,))), t)
        groups.setdefault(key, []).append(v)
    blocks = []
    for (members, t), vs in groups.items():
        if sorted(vs) != list(members):
            continue
        if all(runs[v][3] == tuple(m for m in members if m != v) for v in vs):
            blocks.append((members, t))
    blocks.sort(key=lambda b: b[0][0])
    in_block = np.zeros(src.shape[0], dtype=bool)            # by dst-sorted position
    for members, _ in blocks:
        for v in members:
            rb, ln = runs[v][0], runs[v][1]
            in_block[rb:rb + ln] = True
    return blocks, runs, order, ptr, in_block


def check_view(wb, src, dst, etype, n_nodes):
    dev = torch.device("cuda")
    plan = wb.GraphPlan(src.to(dev), dst.to(dev), n_nodes)
    es = plan.sort_edge_ints(etype.to(dev))
    view = plan.clique_view(es)
    blocks, runs, order, ptr, in_block = np_blocks(src.numpy(), dst.numpy(), etype.numpy(), n_nodes)
    assert view.n_blocks == len(blocks)
    if not blocks:
        return view, blocks
    bp = view.blk_ptr[: view.n_blocks + 1].cpu().numpy()
    bn = view.blk_nodes[: view.n_members].cpu().numpy()
    bt = view.blk_type[: view.n_blocks].cpu().numpy()
    nb = view.node_blk.cpu().numpy()
    nr = view.node_run.cpu().numpy()
    nk = view.node_rank.cpu().numpy()
    soff = view.blk_soff[: view.n_blocks + 1].cpu().numpy()
    exp_nb = -np.ones(n_nodes, dtype=np.int64)
    so = 0
    for i, (members, t) in enumerate(blocks):
        assert tuple(bn[bp[i]:bp[i + 1]]) == members
        assert bt[i] == t
        assert soff[i] == so
        so += len(members) * ((len(members) + 3) // 4 * 4)
        for r, v in enumerate(members):
            exp_nb[v] = i
            assert nr[v] == runs[v][0] and nk[v] == r
    assert soff[len(blocks)] == so == view.s_elems
    assert np.array_equal(nb, exp_nb)
    assert view.max_block == max(len(m) for m, _ in blocks)
    # residual CSR, destination side: the dst-sorted positions outside the block runs, in order
    r_in_ptr, r_in_src, r_in_pos, r_out_ptr, r_out_dst, r_out_pos = [t.cpu().numpy() for t in view.r]
    keep = np.nonzero(~in_block)[0]
    assert view.n_res_edges == keep.shape[0]
    assert np.array_equal(r_in_pos, keep)
    assert np.array_equal(r_in_src, src.numpy()[order][keep])
    dst_sorted = dst.numpy()[order]
    assert np.array_equal(r_in_ptr, np.concatenate([[0], np.cumsum(np.bincount(dst_sorted[keep], minlength=n_nodes))]))
    # source side: stable by (src, original edge id), restricted to the residual edges
    eid_keep = order[keep]
    pos_of_eid = np.empty(src.shape[0], dtype=np.int64)
    pos_of_eid[order] = np.arange(src.shape[0])
    eids = np.sort(eid_keep)
    o2 = eids[np.argsort(src.numpy()[eids], kind="stable")]
    assert np.array_equal(r_out_dst, dst.numpy()[o2])
    assert np.array_equal(r_out_pos, pos_of_eid[o2])
    assert np.array_equal(r_out_ptr, np.concatenate([[0], np.cumsum(np.bincount(src.numpy()[o2], minlength=n_nodes))]))
    return view, blocks


@pytest.mark.parametrize("n_j,n_m,batch,progress", [(10, 10, 3, 0.0), (15, 15, 2, 0.5), (30, 20, 2, 1.0),
                                                    (100, 20, 1, 0.3), (6, 4, 3, 0.0)])
def test_clique_detection_bit_exact(wb, n_j, n_m, batch, progress):
    from wheatley_b200.workload import jssp_batch
    hb = jssp_batch(n_j, n_m, batch, seed0=3, progress=progress)
    view, blocks = check_view(wb, hb["src"], hb["dst"], hb["etype"], hb["n_nodes"])
    if n_j >= 8:
        assert len(blocks) == batch * n_m and all(len(m) == n_j for m, _ in blocks)
    else:
        assert len(blocks) == 0          # 6-node cliques stay on the sparse kernels (MIN_CLIQUE = 8)


def test_clique_detection_rejects_near_cliques(wb):
    """A clique with one edge missing, one with a foreign type inside, one listed in descending order and a
    true one: only the true one may become a block; everything else must stay in the residual CSR."""
    rng = np.random.default_rng(0)
    S, D, T = [], [], []
    def clique(nodes, t=5, drop=None, bad_type=None, descending=False):
        for c in nodes:
            rs = [r for r in nodes if r != c]
            if descending:
                rs = rs[::-1]
            for r in rs:
                if drop == (r, c):
                    continue
                S.append(r); D.append(c); T.append(7 if bad_type == (r, c) else t)
    clique(list(range(0, 12)))
    clique(list(range(12, 24)), drop=(14, 20))
    clique(list(range(24, 36)), bad_type=(25, 30))
    clique(list(range(36, 48)), descending=True)
    clique(list(range(48, 60)), t=6)
    n = 60
    extra = rng.integers(0, n, (2, 100))
    S += list(extra[0]); D += list(extra[1]); T += [1] * 100
    S += list(range(n)); D += list(range(n)); T += [0] * n
    src, dst, et = (torch.tensor(x, dtype=torch.int64) for x in (S, D, T))
    view, blocks = check_view(wb, src, dst, et, n)
    got = sorted(m[0] for m, _ in blocks)
    assert 12 not in got and 24 not in got and 36 not in got


def grid(t, q):
    """Round to multiples of 1/q.  With node features / edge-type embeddings on a 1/8 grid in [-2, 2] and
    the weights feeding the pre-activation on a 1/32 grid, every pre-activation z = Pi[u] + Pj[v] + Pe + b is
    computed EXACTLY in fp32 (all partial sums fit 24 bits), so the kernel and the fp64 oracle take the same
    side of the leaky_relu kink on every element.  Without this, roughly one element in 1e7 lands within
    fp32 rounding of z = 0, flips its derivative and moves single gradient entries by up to 1e-2 relative -
    a property of any fp32 implementation of the layer, not of these kernels."""
    return torch.round(t.clamp(-2, 2) * q) / q


def _egat_case(wb, hb, H, D, gen, no_grad=False, typed=True):
    from oracle import gat_oracle
    Fin = 24
    N = hb["n_nodes"]
    src, dst, etype = hb["src"], hb["dst"], hb["etype"]
    ref = gat_oracle.OracleEGATConv(Fin, Fin, D, D, H).double()
    with torch.no_grad():
        for k, p in ref.named_parameters():
            p.add_(0.1 * torch.randn(p.shape, generator=gen, dtype=p.dtype))
            if k.split(".")[0] in ("fc_ni", "fc_nj", "fc_fij", "bias"):
                p.copy_(grid(p, 32))
    layer = wb.EGATConv(Fin, Fin, D, D, H).cuda()
    layer.load_state_dict({k: v.detach().float() for k, v in ref.state_dict().items()})
    x = grid(torch.randn(N, Fin, generator=gen, dtype=torch.float64), 8)
    Tn = int(etype.max()) + 1
    table = grid(torch.randn(Tn, Fin, generator=gen, dtype=torch.float64), 8)
    R = torch.randn(N, H, D, generator=gen, dtype=torch.float64)
    xr = x.clone().requires_grad_(True)
    tr = table.clone().requires_grad_(True)
    h_ref, _, a_ref = ref(src, dst, xr, tr[etype], get_attention=True)
    (h_ref * R).sum().backward()
    xg = x.float().cuda().requires_grad_(True)
    tg = table.float().cuda().requires_grad_(True)
    plan = wb.GraphPlan(src.cuda(), dst.cuda(), N)
    h, _, a = layer(plan, xg, wb.TypedEdgeFeats(etype.cuda(), tg), get_attention=True)
    (h * R.float().cuda()).sum().backward()
    return dict(h=(h, h_ref), a=(a, a_ref), gx=(xg.grad, xr.grad), gt=(tg.grad, tr.grad),
                params=[(k, p.grad, pr.grad) for (k, p), (_, pr) in
                        zip(layer.named_parameters(), ref.named_parameters())]), plan


@pytest.mark.parametrize("n_j,n_m,batch,progress", [(10, 10, 3, 0.5), (15, 15, 2, 0.0), (30, 20, 1, 0.5),
                                                    (20, 5, 2, 1.0), (9, 3, 4, 0.0),
                                                    (100, 20, 1, 0.0),     # the bench regime: 100-member blocks
                                                    (128, 2, 1, 0.3), (77, 3, 2, 1.0)])
def test_egat_clique_path_matches_oracle(wb, n_j, n_m, batch, progress):
    from wheatley_b200.workload import jssp_batch
    hb = jssp_batch(n_j, n_m, batch, seed0=11, progress=progress)
    gen = torch.Generator().manual_seed(n_j * 100 + n_m)
    res, plan = _egat_case(wb, hb, 4, 64, gen)
    views = list(plan._clique_views.values())
    assert views and views[0].n_blocks == batch * n_m, "the dense clique path was not taken"
    for k in ("h", "a", "gx", "gt"):
        assert_close(res[k][0], res[k][1], k)
    for k, g, gr in res["params"]:
        assert_close(g, gr, "grad " + k)


def test_clique_path_agrees_with_sparse_path(wb, monkeypatch):
    """Same inputs through the dense clique kernels and through the sparse kernels only."""
    from wheatley_b200.workload import jssp_batch
    hb = jssp_batch(12, 6, 3, seed0=5, progress=0.4)
    outs = []
    for flag in ("1", "0"):
        monkeypatch.setenv("WHEATLEY_B200_CLIQUES", flag)
        gen = torch.Generator().manual_seed(99)
        torch.manual_seed(7)                      # identical layer initialisation in both runs
        res, plan = _egat_case(wb, hb, 4, 64, gen)
        outs.append(res)
        assert (list(plan._clique_views.values())[0].n_blocks > 0) == (flag == "1")
    for k in ("h", "a", "gx", "gt"):
        assert_close(outs[0][k][0], outs[1][k][0], k + " dense vs sparse", rel=2e-5)
    for (k, g0, _), (_, g1, _) in zip(outs[0]["params"], outs[1]["params"]):
        assert_close(g0, g1, "grad " + k + " dense vs sparse", rel=2e-5)


@pytest.mark.parametrize("efmode", ["typed", "none"])
def test_gatv2_clique_path_matches_oracle(wb, efmode):
    """PyG flavour: V aliases x_l, slope 0.2, softmax +1e-16, output bias."""
    from oracle import gat_oracle
    from wheatley_b200.workload import jssp_batch
    hb = jssp_batch(14, 5, 2, seed0=2, progress=0.3)
    N, src, dst, etype = hb["n_nodes"], hb["src"], hb["dst"], hb["etype"]
    gen = torch.Generator().manual_seed(4)
    Fin, D, H = 32, 64, 2
    ref = gat_oracle.OracleGATv2Conv(Fin, D, H, edge_dim=Fin).double()
    with torch.no_grad():
        for k, p in ref.named_parameters():
            p.add_(0.1 * torch.randn(p.shape, generator=gen, dtype=p.dtype))
            if k.split(".")[0] in ("lin_l", "lin_r", "lin_edge"):
                p.copy_(grid(p, 32))
    layer = wb.GATv2Conv(Fin, D, H, edge_dim=Fin).cuda()
    layer.load_state_dict({k: v.detach().float() for k, v in ref.state_dict().items()})
    x = grid(torch.randn(N, Fin, generator=gen, dtype=torch.float64), 8)
    table = grid(torch.randn(int(etype.max()) + 1, Fin, generator=gen, dtype=torch.float64), 8)
    R = torch.randn(N, H * D, generator=gen, dtype=torch.float64)
    xr = x.clone().requires_grad_(True)
    tr = table.clone().requires_grad_(True)
    ei = torch.stack([src, dst])
    out_ref = ref(xr, ei, tr[etype] if efmode == "typed" else None)
    (out_ref * R).sum().backward()
    xg = x.float().cuda().requires_grad_(True)
    tg = table.float().cuda().requires_grad_(True)
    plan = wb.GraphPlan(src.cuda(), dst.cuda(), N)
    out = layer(xg, plan, wb.TypedEdgeFeats(etype.cuda(), tg) if efmode == "typed" else None)
    (out * R.float().cuda()).sum().backward()
    nb = list(plan._clique_views.values())[0].n_blocks
    # untyped: a neighbouring non-clique edge can extend a node's ascending run, which (correctly) rejects its block
    assert nb == 2 * 5 if efmode == "typed" else nb >= 1
    assert_close(out, out_ref, "out")
    assert_close(xg.grad, xr.grad, "grad x")
    if efmode == "typed":
        assert_close(tg.grad, tr.grad, "grad table")
    refp = dict(ref.named_parameters())
    for k, p in layer.named_parameters():
        if p.grad is not None and refp[k].grad is not None:
            assert_close(p.grad, refp[k].grad, "grad " + k)
