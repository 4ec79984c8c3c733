"""Whole-trunk passes: the L (projections -> fused attention -> GELU MLP (+ residual)) layers of
jssp/models/gnn_dgl.py:366-399 / psp/models/gnn_flat.py:170-253 behind ONE autograd node that makes one C-ABI call
per pass (``b200gat_trunk_fwd`` / ``b200gat_trunk_bwd``, csrc/trunk.cu).

Why: the eager PPO step of the layer-by-layer mirror is host-bound - ~300 kernel launches issued from Python through
~40 autograd nodes cost 7.4 ms per step against 4.6 ms of device work.  The same launches issued from C cost ~1 ms.
The per-layer modules (``EGATConv``, ``GraphConv``, ``MLP``) stay the drop-in boundary and run the same kernels; the
trunks (``GnnDGL``, ``GnnFlat``) take this path whenever the configuration is covered (``trunk_applicable``)."""
import ctypes as C
import os

import torch

from . import _lib
from ._lib import eager_only
from .dense import tf32_path_enabled
from .plan import check_edge_types

MAXL, MAXM = _lib.TRUNK_MAX_LAYERS, _lib.MLP_MAX_LAYERS


class LayerSpec:
    """Parameters of one trunk layer, as tensors (the modules keep owning them)."""

    def __init__(self, proj_w, proj_b, attn, out_bias, table, affine, mlp_w, mlp_b):
        self.proj_w, self.proj_b = list(proj_w), (None if proj_b is None else list(proj_b))
        self.attn, self.out_bias, self.table, self.affine = attn, out_bias, table, affine
        self.mlp_w, self.mlp_b = list(mlp_w), list(mlp_b)

    def flat(self):
        pb = self.proj_b if self.proj_b is not None else [None] * len(self.proj_w)
        return [*self.proj_w, *pb, self.attn, self.out_bias, self.table, self.affine, *self.mlp_w, *self.mlp_b]


def trunk_enabled() -> bool:
    return os.environ.get("WHEATLEY_B200_TRUNK", "fused") == "fused"


def _c(t):
    return None if t is None else (t if t.is_contiguous() else t.contiguous())


def _fill_layer(L, cfg, tensors, etype_sorted, attr_sorted, edge_mode):
    """tensors: the layer's slice of the flat parameter list (contiguous fp32)."""
    n_proj, n_mlp = cfg["n_proj"], cfg["n_mlp"]
    it = iter(tensors)
    pw = [next(it) for _ in range(n_proj)]
    pb = [next(it) for _ in range(n_proj)]
    attn, out_bias, table, affine = next(it), next(it), next(it), next(it)
    mw = [next(it) for _ in range(n_mlp)]
    mb = [next(it) for _ in range(n_mlp)]
    L.n_proj = n_proj
    for p in range(n_proj):
        L.Wp[p] = pw[p].data_ptr()
        L.bp[p] = _lib.ptr(pb[p])
    L.attn = _lib.Attn(_lib.FLAVOUR_ADDITIVE, cfg["heads"], cfg["dim"], cfg["slope"], cfg["eps"], attn.data_ptr(),
                       None, _lib.ptr(out_bias))
    L.ef = _lib.EdgeFeats(edge_mode, 0 if table is None else table.shape[0], _lib.ptr(table),
                          _lib.ptr(etype_sorted) if table is not None else None, None,
                          0 if affine is None else affine.shape[0],
                          _lib.ptr(attr_sorted) if affine is not None else None, _lib.ptr(affine))
    L.n_mlp = n_mlp
    L.dims[0] = mw[0].shape[1]
    for i in range(n_mlp):
        L.dims[i + 1] = mw[i].shape[0]
        L.W[i] = mw[i].data_ptr()
        L.b[i] = _lib.ptr(mb[i])
    return pw, pb, attn, out_bias, table, affine, mw, mb


def setup_trunk(t, view, cfg, etype_sorted, attr_sorted, x0, flat, need):
    """Fill the ``b200gat_trunk`` struct ``t`` (a ``_lib.Trunk``, possibly embedded in a larger struct) and allocate
    what the forward writes.  Returns ``(flat_c, keep, y)``: the contiguous parameter tensors, every tensor whose
    storage ``t`` points into (keep them alive as long as ``t`` is used) and the ``[L, N, in_dim]`` layer outputs."""
    nl, n_proj, n_mlp, H, D = cfg["n_layers"], cfg["n_proj"], cfg["n_mlp"], cfg["heads"], cfg["dim"]
    per = 2 * n_proj + 4 + 2 * n_mlp
    assert len(flat) == nl * per
    dev = x0.device
    flat_c = [_c(None if f is None else (f.reshape(-1) if i % per == 2 * n_proj else f)) for i, f in enumerate(flat)]
    N, E, HD, in_dim = x0.shape[0], int(view.c.n_edges), H * D, x0.shape[1]
    t.n_layers, t.in_dim, t.residual, t.n_nodes = nl, in_dim, int(cfg["residual"]), N
    t.graph = C.pointer(view.c)
    t.x0 = x0.data_ptr()
    for l in range(nl):
        filled = _fill_layer(t.layer[l], cfg, flat_c[l * per:(l + 1) * per], etype_sorted, attr_sorted,
                             cfg["edge_mode"])
        if filled[4] is not None:            # IndexError like nn.Embedding instead of an out-of-bounds gather
            check_edge_types(etype_sorted, filled[4].shape[0])
    if not _lib.lib.b200gat_trunk_supported(C.byref(t)):
        raise _lib.B200GatError("trunk configuration not covered by b200gat_trunk_fwd (use the layer modules)")
    f32 = dict(dtype=torch.float32, device=dev)
    wcat = torch.empty(int(_lib.lib.b200gat_trunk_wcat_elems(C.byref(t))), **f32)
    t.wcat = wcat.data_ptr()
    dense = bool(_lib.lib.b200gat_conv_uses_cliques(t.graph, C.byref(t.layer[0].attn), C.byref(t.layer[0].ef)))
    s_el = H * getattr(view, "s_elems", 0) if dense else 0
    # saved activations: one allocation per kind for all layers (every layer's slice is read by the backward);
    # without autograd every layer re-uses the first slice
    nk = nl if need else 1
    hid = [t.layer[0].dims[i + 1] for i in range(n_mlp - 1)]
    P = torch.empty((nk, n_proj, N, HD), **f32)
    conv_out = torch.empty((nk, N, HD), **f32)
    s = torch.empty((nk, max(E, 1), H), **f32)
    ml = torch.empty((nk, 2, N, H), **f32)
    agg = torch.empty((nk, N, HD), **f32) if (need or dense) else None
    s_dense = torch.empty((nk, max(s_el, 4)), **f32) if dense else None
    hs = [torch.empty((nk, N, w), **f32) for w in hid]
    zs = [torch.empty((nk, N, w), **f32) for w in hid] if need else []
    y = torch.empty((nl, N, in_dim), **f32)
    ptr = lambda a, k: a.data_ptr() + k * a.stride(0) * 4                          # noqa: E731
    for l in range(nl):
        L, k = t.layer[l], (l if need else 0)
        L.P, L.conv_out, L.s_save = ptr(P, k), ptr(conv_out, k), ptr(s, k)
        L.m_save = ptr(ml, k)
        L.l_save = L.m_save + ml.stride(1) * 4
        L.agg_save = None if agg is None else ptr(agg, k)
        L.s_dense = None if s_dense is None else ptr(s_dense, k)
        for i in range(n_mlp - 1):
            L.h[i] = ptr(hs[i], k)
            L.z[i] = ptr(zs[i], k) if zs else None
        L.y = ptr(y, l)
    keep = [x0, *flat_c, wcat, P, conv_out, s, ml, agg, s_dense, *hs, *zs, y, etype_sorted, attr_sorted]
    return flat_c, keep, y


def setup_trunk_grads(g, cfg, flat_c, shapes, dev):
    """Fill the per-layer parameter-gradient pointers of ``g`` (a ``_lib.TrunkGrads``) and return the gradient
    tensors in the order of the flat parameter list.  A handful of allocations for all layers: the concatenated
    projection gradients, one zero-filled buffer for the atomically accumulated ones, one stacked tensor per MLP
    weight / bias position."""
    nl, n_proj, n_mlp, H, D = cfg["n_layers"], cfg["n_proj"], cfg["n_mlp"], cfg["heads"], cfg["dim"]
    per = 2 * n_proj + 4 + 2 * n_mlp
    HD = H * D
    in_dim = flat_c[0].shape[1]
    f32 = dict(dtype=torch.float32, device=dev)
    sizes = []
    for l in range(nl):
        attn, _ob, table, affine = flat_c[l * per + 2 * n_proj: l * per + 2 * n_proj + 4]
        sizes += [(attn.numel() + 3) // 4 * 4, 0 if table is None else (table.numel() + 3) // 4 * 4,
                  0 if affine is None else (affine.numel() + 3) // 4 * 4]
    zbuf = torch.zeros(sum(sizes), **f32)
    dWp = torch.empty((nl, n_proj * HD, in_dim), **f32)
    has_pb = flat_c[n_proj] is not None
    dbp = torch.empty((nl, n_proj * HD), **f32) if has_pb else None
    has_ob = flat_c[2 * n_proj + 1] is not None
    g_ob = torch.empty((nl, HD), **f32) if has_ob else None
    mw = [flat_c[2 * n_proj + 4 + i] for i in range(n_mlp)]
    dWs = [torch.empty((nl,) + tuple(w.shape), **f32) for w in mw]                 # layers share their MLP shapes
    dbs = [torch.empty((nl, w.shape[0]), **f32) if flat_c[2 * n_proj + 4 + n_mlp + i] is not None else None
           for i, w in enumerate(mw)]
    grads = [None] * (nl * per)
    off = 0
    zp = zbuf.data_ptr()
    for l in range(nl):
        G = g.layer[l]
        base = l * per
        attn, _ob, table, affine = flat_c[base + 2 * n_proj: base + 2 * n_proj + 4]
        G.dWp = dWp.data_ptr() + l * dWp.stride(0) * 4
        G.dbp = None if dbp is None else dbp.data_ptr() + l * dbp.stride(0) * 4
        dWp_l = dWp[l]
        for p in range(n_proj):
            grads[base + p] = dWp_l[p * HD:(p + 1) * HD]
            if has_pb:
                grads[base + n_proj + p] = dbp[l, p * HD:(p + 1) * HD]
        G.g_attn = zp + 4 * off
        grads[base + 2 * n_proj] = zbuf[off: off + attn.numel()].view(shapes[base + 2 * n_proj])
        off += sizes[3 * l]
        if has_ob:
            G.g_out_bias = g_ob.data_ptr() + l * HD * 4
            grads[base + 2 * n_proj + 1] = g_ob[l]
        if table is not None:
            G.g_table = zp + 4 * off
            grads[base + 2 * n_proj + 2] = zbuf[off: off + table.numel()].view(table.shape)
        off += sizes[3 * l + 1]
        if affine is not None:
            G.g_affine = zp + 4 * off
            grads[base + 2 * n_proj + 3] = zbuf[off: off + affine.numel()].view(affine.shape)
        off += sizes[3 * l + 2]
        for i in range(n_mlp):
            G.dW[i] = dWs[i].data_ptr() + l * dWs[i].stride(0) * 4
            grads[base + 2 * n_proj + 4 + i] = dWs[i][l]
            if dbs[i] is not None:
                G.db[i] = dbs[i].data_ptr() + l * dbs[i].stride(0) * 4
                grads[base + 2 * n_proj + 4 + n_mlp + i] = dbs[i][l]
    return grads


class _TrunkFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, view, cfg, etype_sorted, attr_sorted, x0, *flat):
        x0 = _c(x0)
        need = any(ctx.needs_input_grad)
        t = _lib.Trunk()
        flat_c, keep, y = setup_trunk(t, view, cfg, etype_sorted, attr_sorted, x0, flat, need)
        with _lib.device_guard(x0.device):
            _lib.check(_lib.lib.b200gat_trunk_fwd(C.byref(t), _lib.stream_ptr()), "b200gat_trunk_fwd")
        if need:
            ctx.t, ctx.view, ctx.cfg = t, view, cfg
            ctx.n_flat = len(flat_c)
            ctx.save_for_backward(*keep)          # raw pointers in ctx.t stay valid while these are alive
            ctx.shapes = [None if f is None else f.shape for f in flat]
            ctx.set_materialize_grads(False)
        return tuple(y.unbind(0))

    @staticmethod
    def backward(ctx, *g_ys):
        t, cfg = ctx.t, ctx.cfg
        nl = cfg["n_layers"]
        saved = ctx.saved_tensors
        x0, flat_c = saved[0], list(saved[1: 1 + ctx.n_flat])
        if all(g is None for g in g_ys):
            return (None,) * (5 + len(flat_c))
        dev = x0.device
        f32 = dict(dtype=torch.float32, device=dev)
        g = _lib.TrunkGrads()
        g_ys = [None if gy is None else _c(gy) for gy in g_ys]
        if g_ys[-1] is None:
            g_ys[-1] = torch.zeros(tuple(x0.shape), **f32)
        grads = setup_trunk_grads(g, cfg, flat_c, ctx.shapes, dev)
        for l in range(nl):
            g.layer[l].g_y = _lib.ptr(g_ys[l])
        g_x0 = torch.empty_like(x0) if ctx.needs_input_grad[4] else None
        g.g_x0 = _lib.ptr(g_x0)
        ws_elems = int(_lib.lib.b200gat_trunk_bwd_workspace_elems(C.byref(t)))
        ws = torch.empty(max(ws_elems, 4), **f32)
        g.workspace, g.workspace_elems = ws.data_ptr(), ws_elems
        with _lib.device_guard(dev):
            _lib.check(_lib.lib.b200gat_trunk_bwd(C.byref(t), C.byref(g), _lib.stream_ptr()), "b200gat_trunk_bwd")
        need = ctx.needs_input_grad[5:]
        grads = [gr if nd else None for gr, nd in zip(grads, need)]
        return (None, None, None, None, g_x0, *grads)


def trunk_cfg(specs, heads, dim, slope, eps, residual):
    s0 = specs[0]
    mode = _lib.EDGE_NONE if s0.table is None else (_lib.EDGE_TABLE if s0.affine is None else _lib.EDGE_AFFINE)
    return dict(n_layers=len(specs), n_proj=len(s0.proj_w), n_mlp=len(s0.mlp_w), heads=int(heads), dim=int(dim),
                slope=float(slope), eps=float(eps), residual=bool(residual), edge_mode=mode)


@eager_only
def fused_trunk(view, etype_sorted, attr_sorted, x0, specs, *, heads, dim, slope, eps, residual):
    """Outputs of the L layers, ``[N, in_dim]`` each.  ``view``: the plan / clique view the layers' edge mode
    selects (``EdgeFeatures.graph_view``); ``etype_sorted`` / ``attr_sorted``: per-edge type ids / attributes in the
    plan's destination-sorted order (shared by all layers)."""
    if not x0.is_cuda:
        raise _lib.B200GatError("wheatley_b200 kernels need CUDA tensors (no CPU fallback)")
    cfg = trunk_cfg(specs, heads, dim, slope, eps, residual)
    flat = []
    for sp in specs:
        flat += sp.flat()
    return list(_TrunkFn.apply(view, cfg, etype_sorted, attr_sorted, x0, *flat))


def trunk_applicable(x0, specs, heads, dim) -> bool:
    """Static part of the coverage test (the C side re-checks the GEMM shapes)."""
    if not (trunk_enabled() and tf32_path_enabled() and x0.is_cuda and x0.dtype == torch.float32 and x0.dim() == 2
            and x0.shape[0] > 0):
        return False
    if not specs or len(specs) > MAXL:
        return False
    HD, in_dim = heads * dim, x0.shape[1]
    if HD % 128 or in_dim % 32:
        return False
    s0 = specs[0]
    for sp in specs:
        if (len(sp.proj_w) != len(s0.proj_w) or len(sp.mlp_w) != len(s0.mlp_w) or len(sp.mlp_w) > MAXM
                or (sp.proj_b is None) != (s0.proj_b is None) or (sp.table is None) != (s0.table is None)
                or (sp.affine is None) != (s0.affine is None) or (sp.out_bias is None) != (s0.out_bias is None)):
            return False
        if any(tuple(w.shape) != (HD, in_dim) for w in sp.proj_w):
            return False
        ws = [w.shape for w in sp.mlp_w]
        if ws[0][1] != HD or ws[-1][0] != in_dim or any(w[0] % 4 or w[1] % 4 for w in ws):
            return False
        if [tuple(w) for w in ws] != [tuple(w.shape) for w in s0.mlp_w]:
            return False
        if sp.affine is not None and sp.affine.shape[0] > _lib.MAX_ATTR:
            return False
    return True
