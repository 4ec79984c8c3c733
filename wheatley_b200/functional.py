"""autograd.Function wrappers around the C-ABI kernels (the only place the extension is
called from).  No PyTorch fallback exists: CPU tensors raise."""
import torch

from . import _lib
from ._lib import eager_only
from .plan import GraphPlan, check_edge_types


class EdgeFeatures:
    """Factored edge features handed to the fused layer.

    ``mode='table'``: ``Pe[e] = table[etype[e]]`` with ``table [T, H*D]`` already projected
    (the reference computes ``fc_fij(Embedding(type))`` per edge, jssp/models/gnn_dgl.py:364 +
    generic/eegatconv.py:262; projecting the T-row table instead is the same map).
    ``mode='dense'``: ``Pe [E, H*D]`` in the original edge order.
    """

    def __init__(self, mode, table=None, etype_sorted=None, dense=None, attr_sorted=None, affine=None):
        self.mode, self.table, self.etype_sorted, self.dense = mode, table, etype_sorted, dense
        self.attr_sorted, self.affine = attr_sorted, affine     # 'affine' mode: [E,K] sorted, [K,H*D]

    def graph_view(self, plan):
        """The graph descriptor the kernels get: with clique blocks for the table / untyped modes
        (a block needs one edge type), the plain plan otherwise."""
        if self.mode == _lib.EDGE_TABLE:
            return plan.clique_view(self.etype_sorted)
        if self.mode == _lib.EDGE_NONE:
            return plan.clique_view(None)
        return plan

    @classmethod
    def none(cls):
        return cls(_lib.EDGE_NONE)


@eager_only
def project_edge_feats(plan: GraphPlan, efeats, weight, pre_bias=None) -> EdgeFeatures:
    """Edge term of the logit, ``Pe = efeats @ weight.T (+ pre_bias)``, in the cheapest exact form:
    typed/affine factors are projected as a T-row table (+ a K-row affine map) and gathered
    inside the kernel; a dense ``[E, F]`` tensor is projected with one GEMM."""
    from .edge_feats import TypedEdgeFeats
    if efeats is None:
        return EdgeFeatures.none()
    if isinstance(efeats, TypedEdgeFeats):
        table = efeats.tables.get(id(weight)) if getattr(efeats, "tables", None) else None
        if table is None or efeats.attrs is not None:
            table = torch.nn.functional.linear(efeats.weight, weight, pre_bias)  # [T, H*D]
        etype_sorted = plan.sort_edge_ints(efeats.etype)
        if efeats.attrs is not None and efeats.attrs.shape[1] <= _lib.MAX_ATTR:
            affine = (weight @ efeats.attr_weight).t()                            # [K, H*D]
            return EdgeFeatures(_lib.EDGE_AFFINE, table=table, etype_sorted=etype_sorted,
                                attr_sorted=plan.sort_edge_rows(efeats.attrs), affine=affine)
        if efeats.attrs is None:
            return EdgeFeatures(_lib.EDGE_TABLE, table=table, etype_sorted=etype_sorted)
        efeats = efeats.dense()
    return EdgeFeatures(_lib.EDGE_DENSE, dense=torch.nn.functional.linear(efeats, weight, pre_bias))


def _f32c(t):
    if t is None:
        return None
    if t.dtype != torch.float32:
        raise _lib.B200GatError(f"fp32 tensors required, got {t.dtype}")
    return t.contiguous()


class _FusedGATConv(torch.autograd.Function):
    """out[v] = sum_{e->v} softmax_v(s_e) * V[src_e],  s_e by flavour (include/b200gat.h)."""

    @staticmethod
    def forward(ctx, plan: GraphPlan, view, cfg: dict, Pi, Pj, V, attn, attn_bias, out_bias,
                e_table, e_dense, etype_sorted, e_affine, attr_sorted):
        if not Pi.is_cuda:
            raise _lib.B200GatError("wheatley_b200 kernels need CUDA tensors (no CPU fallback)")
        H, D = cfg["heads"], cfg["dim"]
        N, E, HD = plan.n_nodes, plan.n_edges, H * D
        Pi, Pj = _f32c(Pi), _f32c(Pj)
        v_alias = V is None
        Vt = Pi if v_alias else _f32c(V)
        attn_c = _f32c(attn.reshape(-1))
        attn_bias_c = _f32c(attn_bias)
        out_bias_c = _f32c(out_bias)
        e_table_c, e_dense_c, e_affine_c = _f32c(e_table), _f32c(e_dense), _f32c(e_affine)
        mode = _lib.EDGE_TABLE if e_table is not None else (_lib.EDGE_DENSE if e_dense is not None
                                                           else _lib.EDGE_NONE)
        if e_affine is not None:
            mode = _lib.EDGE_AFFINE
        n_attr = 0 if e_affine_c is None else e_affine_c.shape[0]
        if e_table_c is not None:           # IndexError like nn.Embedding, instead of an out-of-bounds gather
            check_edge_types(etype_sorted, e_table_c.shape[0])
        assert Pi.shape == (N, HD) and Pj.shape == (N, HD) and Vt.shape == (N, HD)
        out = torch.empty((N, HD), dtype=torch.float32, device=Pi.device)
        s = torch.empty((E, H), dtype=torch.float32, device=Pi.device)
        m = torch.empty((N, H), dtype=torch.float32, device=Pi.device)
        l = torch.empty((N, H), dtype=torch.float32, device=Pi.device)
        need_grad = any(ctx.needs_input_grad)
        agg = torch.empty((N, HD), dtype=torch.float32, device=Pi.device) if need_grad else None
        a = _lib.Attn(cfg["flavour"], H, D, cfg["slope"], cfg["eps"], _lib.ptr(attn_c),
                      _lib.ptr(attn_bias_c), _lib.ptr(out_bias_c))
        ef = _lib.EdgeFeats(mode, 0 if e_table_c is None else e_table_c.shape[0],
                            _lib.ptr(e_table_c), _lib.ptr(etype_sorted), _lib.ptr(e_dense_c),
                            n_attr, _lib.ptr(attr_sorted), _lib.ptr(e_affine_c))
        dense = bool(_lib.lib.b200gat_conv_uses_cliques(view.c, a, ef))
        s_dense = torch.empty(H * view.s_elems, dtype=torch.float32, device=Pi.device) if dense else None
        if dense and agg is None:           # the dense backward needs it; the forward writes it anyway
            agg = torch.empty((N, HD), dtype=torch.float32, device=Pi.device)
        with _lib.device_guard(Pi.device):
            _lib.check(_lib.lib.b200gat_conv_fwd(view.c, a, ef, _lib.ptr(Pi), _lib.ptr(Pj), _lib.ptr(Vt),
                                                 _lib.ptr(out), _lib.ptr(s), _lib.ptr(m), _lib.ptr(l),
                                                 _lib.ptr(agg), _lib.ptr(s_dense), _lib.stream_ptr()),
                       "b200gat_conv_fwd")
        # no zero-filled stand-ins for the gradients of the non-differentiable outputs (s, m, l, s_dense are
        # 2 x 26 MB per layer at 8 x Taillard 100x20: materialising them was 4 memsets per layer)
        ctx.set_materialize_grads(False)
        ctx.plan, ctx.view, ctx.cfg, ctx.mode, ctx.v_alias = plan, view, cfg, mode, v_alias
        ctx.dense = dense
        ctx.etype_sorted, ctx.attr_sorted = etype_sorted, attr_sorted
        ctx.has_out_bias = out_bias is not None
        ctx.attn_shape = attn.shape
        ctx.save_for_backward(Pi, Pj, Vt, attn_c, attn_bias_c, e_table_c, e_dense_c, e_affine_c, s, m, l, agg,
                              s_dense)
        if s_dense is None:
            s_dense = torch.empty(0, dtype=torch.float32, device=Pi.device)
        ctx.mark_non_differentiable(s, m, l, s_dense)
        return out, s, m, l, s_dense

    @staticmethod
    def backward(ctx, g_out, _gs, _gm, _gl, _gsd):
        (Pi, Pj, Vt, attn_c, attn_bias_c, e_table_c, e_dense_c, e_affine_c, s, m, l, agg,
         s_dense) = ctx.saved_tensors
        plan, view, cfg = ctx.plan, ctx.view, ctx.cfg
        H, D = cfg["heads"], cfg["dim"]
        if g_out is None:
            return (None,) * 14
        g_out = _f32c(g_out)
        # one [3, N, HD] buffer: the projection backward (dense._Proj3) reads gPi | gPj | gV as one operand
        g3 = torch.empty((2 if ctx.v_alias else 3,) + tuple(Pi.shape), dtype=Pi.dtype, device=Pi.device)
        gPi, gPj = g3[0], g3[1]
        gV = gPi if ctx.v_alias else g3[2]
        # the small gradients the kernels accumulate into with atomics: one zero-filled buffer, carved up
        # (16-byte aligned pieces)
        acc = [attn_c, attn_bias_c, e_table_c, e_affine_c]
        sizes = [0 if t is None else (t.numel() + 3) // 4 * 4 for t in acc]
        zbuf = torch.zeros(sum(sizes), dtype=attn_c.dtype, device=attn_c.device)
        offs = [sum(sizes[:i]) for i in range(4)]
        g_attn, g_attn_bias, g_table, g_affine = (
            None if t is None else zbuf[o:o + t.numel()].view(t.shape) for t, o in zip(acc, offs))
        g_dense = torch.empty_like(e_dense_c) if e_dense_c is not None else None
        n_gs = 2 * s.numel()
        if ctx.dense:       # + gs_dense [H*s_elems] + gsum_u / gsum_v [N*H]   (include/b200gat.h)
            n_gs += ((H * view.s_elems + 3) // 4) * 4 + 2 * plan.n_nodes * H
        gs = torch.empty(n_gs, dtype=s.dtype, device=s.device)
        a = _lib.Attn(cfg["flavour"], H, D, cfg["slope"], cfg["eps"], _lib.ptr(attn_c),
                      _lib.ptr(attn_bias_c), None)
        ef = _lib.EdgeFeats(ctx.mode, 0 if e_table_c is None else e_table_c.shape[0],
                            _lib.ptr(e_table_c), _lib.ptr(ctx.etype_sorted), _lib.ptr(e_dense_c),
                            0 if e_affine_c is None else e_affine_c.shape[0], _lib.ptr(ctx.attr_sorted),
                            _lib.ptr(e_affine_c))
        with _lib.device_guard(Pi.device):
            _lib.check(_lib.lib.b200gat_conv_bwd(
                view.c, a, ef, _lib.ptr(Pi), _lib.ptr(Pj), _lib.ptr(Vt), _lib.ptr(s), _lib.ptr(m),
                _lib.ptr(l), _lib.ptr(agg), _lib.ptr(s_dense), _lib.ptr(g_out), _lib.ptr(gPi), _lib.ptr(gPj),
                _lib.ptr(gV),
                _lib.ptr(g_attn), _lib.ptr(g_attn_bias), _lib.ptr(g_table), _lib.ptr(g_dense),
                _lib.ptr(g_affine), _lib.ptr(gs), _lib.stream_ptr()), "b200gat_conv_bwd")
        g_bias = g_out.sum(0) if ctx.has_out_bias else None
        return (None, None, None, gPi, gPj, None if ctx.v_alias else gV, g_attn.view(ctx.attn_shape),
                g_attn_bias, g_bias, g_table, g_dense, None, g_affine, None)


@eager_only
def fused_gat_conv(plan, Pi, Pj, V, attn, *, heads, dim, flavour=_lib.FLAVOUR_ADDITIVE, slope=0.2,
                   eps=0.0, attn_bias=None, out_bias=None, edge_feats: EdgeFeatures = None):
    """Returns ``(out [N, H*D], (s, m, l))``; ``V=None`` means V aliases Pi."""
    cfg = dict(heads=int(heads), dim=int(dim), flavour=int(flavour), slope=float(slope), eps=float(eps))
    ef = edge_feats or EdgeFeatures.none()
    e_table = ef.table if ef.mode in (_lib.EDGE_TABLE, _lib.EDGE_AFFINE) else None
    e_dense = ef.dense if ef.mode == _lib.EDGE_DENSE else None
    e_affine = ef.affine if ef.mode == _lib.EDGE_AFFINE else None
    view = ef.graph_view(plan) if flavour == _lib.FLAVOUR_ADDITIVE else plan
    out, s, m, l, s_dense = _FusedGATConv.apply(plan, view, cfg, Pi, Pj, V, attn, attn_bias, out_bias, e_table,
                                                e_dense, ef.etype_sorted if e_table is not None else None,
                                                e_affine, ef.attr_sorted if e_affine is not None else None)
    return out, (s, m, l, s_dense if s_dense.numel() else None, view)


@eager_only
def attention_weights(plan: GraphPlan, saved, heads: int, eps: float) -> torch.Tensor:
    """alpha [E, H] in the original edge order from the logits saved by the forward."""
    s, m, l, s_dense, view = saved
    alpha = torch.empty((plan.n_edges, heads), dtype=torch.float32, device=s.device)
    with _lib.device_guard(s.device):
        _lib.check(_lib.lib.b200gat_attention(view.c, heads, eps, _lib.ptr(s), _lib.ptr(m), _lib.ptr(l),
                                              _lib.ptr(s_dense), _lib.ptr(alpha), _lib.stream_ptr()),
                   "b200gat_attention")
    return alpha


@eager_only
def edge_output(plan: GraphPlan, Pi, Pj, edge_feats: EdgeFeatures, *, heads, dim, slope, attn):
    """f_out [E, H, D] = leaky_relu(Pi[src]+Pj[dst]+Pe) (no autograd: materialised on request)."""
    Pi, Pj = _f32c(Pi.detach()), _f32c(Pj.detach())
    ef = edge_feats or EdgeFeatures.none()
    tbl = _f32c(ef.table.detach()) if ef.mode in (_lib.EDGE_TABLE, _lib.EDGE_AFFINE) else None
    dns = _f32c(ef.dense.detach()) if ef.mode == _lib.EDGE_DENSE else None
    aff = _f32c(ef.affine.detach()) if ef.mode == _lib.EDGE_AFFINE else None
    attn_c = _f32c(attn.detach().reshape(-1))
    a = _lib.Attn(_lib.FLAVOUR_ADDITIVE, heads, dim, slope, 0.0, _lib.ptr(attn_c), None, None)
    efc = _lib.EdgeFeats(ef.mode, 0 if tbl is None else tbl.shape[0], _lib.ptr(tbl),
                         _lib.ptr(ef.etype_sorted), _lib.ptr(dns), 0 if aff is None else aff.shape[0],
                         _lib.ptr(ef.attr_sorted), _lib.ptr(aff))
    f = torch.empty((plan.n_edges, heads, dim), dtype=torch.float32, device=Pi.device)
    with _lib.device_guard(Pi.device):
        _lib.check(_lib.lib.b200gat_edge_out(plan.c, a, efc, _lib.ptr(Pi), _lib.ptr(Pj), _lib.ptr(f),
                                             _lib.stream_ptr()), "b200gat_edge_out")
    return f


class _MaskedHead(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, mask_u8, fill):
        if not logits.is_cuda:
            raise _lib.B200GatError("wheatley_b200 kernels need CUDA tensors (no CPU fallback)")
        logits = _f32c(logits)
        B, n = logits.shape
        logp = torch.empty_like(logits)
        ent = torch.empty(B, dtype=torch.float32, device=logits.device)
        amax = torch.empty(B, dtype=torch.int64, device=logits.device)
        with _lib.device_guard(logits.device):
            _lib.check(_lib.lib.b200gat_mask_head_fwd(
                _lib.ptr(logits), logits.stride(0), _lib.ptr(mask_u8),
                0 if mask_u8 is None else mask_u8.stride(0), B, n, fill, _lib.ptr(logp), _lib.ptr(ent),
                _lib.ptr(amax), _lib.stream_ptr()), "b200gat_mask_head_fwd")
        ctx.save_for_backward(logp, ent, mask_u8)
        ctx.mark_non_differentiable(amax)
        return logp, ent, amax

    @staticmethod
    def backward(ctx, g_logp, g_ent, _g_amax):
        logp, ent, mask_u8 = ctx.saved_tensors
        B, n = logp.shape
        # general path: logp is used through arbitrary downstream ops -> dense softmax backward
        # d logits = g_logp - p * sum(g_logp)  (masked entries get 0), entropy term via the kernel.
        g_logits = torch.empty_like(logp)
        with _lib.device_guard(logp.device):
            _lib.check(_lib.lib.b200gat_mask_head_bwd(
                _lib.ptr(logp), _lib.ptr(mask_u8), 0 if mask_u8 is None else mask_u8.stride(0), None,
                _lib.ptr(ent), None, _lib.ptr(_f32c(g_ent)) if g_ent is not None else None, B, n,
                _lib.ptr(g_logits), _lib.stream_ptr()), "b200gat_mask_head_bwd")
        if g_logp is not None:
            p = logp.exp()
            dense = g_logp - p * g_logp.sum(-1, keepdim=True)
            if mask_u8 is not None:
                dense = dense * mask_u8[:, :n].to(dense.dtype)
            g_logits = g_logits + dense
        return g_logits, None, None


class _MaskedHeadEval(torch.autograd.Function):
    """logprob[action] and entropy with the fused backward (the PPO update call,
    alg/ppo.py:575-581)."""

    @staticmethod
    def forward(ctx, logits, mask_u8, action, fill):
        logits = _f32c(logits)
        B, n = logits.shape
        logp = torch.empty_like(logits)
        ent = torch.empty(B, dtype=torch.float32, device=logits.device)
        with _lib.device_guard(logits.device):
            _lib.check(_lib.lib.b200gat_mask_head_fwd(
                _lib.ptr(logits), logits.stride(0), _lib.ptr(mask_u8),
                0 if mask_u8 is None else mask_u8.stride(0), B, n, fill, _lib.ptr(logp), _lib.ptr(ent),
                None, _lib.stream_ptr()), "b200gat_mask_head_fwd")
        action = action.long().contiguous()
        lp = logp.gather(1, action.view(-1, 1)).squeeze(1)
        ctx.save_for_backward(logp, ent, mask_u8, action)
        return lp, ent

    @staticmethod
    def backward(ctx, g_lp, g_ent):
        logp, ent, mask_u8, action = ctx.saved_tensors
        B, n = logp.shape
        g_logits = torch.empty_like(logp)
        with _lib.device_guard(logp.device):
            _lib.check(_lib.lib.b200gat_mask_head_bwd(
                _lib.ptr(logp), _lib.ptr(mask_u8), 0 if mask_u8 is None else mask_u8.stride(0),
                _lib.ptr(action), _lib.ptr(ent), _lib.ptr(_f32c(g_lp)) if g_lp is not None else None,
                _lib.ptr(_f32c(g_ent)) if g_ent is not None else None, B, n, _lib.ptr(g_logits),
                _lib.stream_ptr()), "b200gat_mask_head_bwd")
        return g_logits, None, None, None


def _mask_u8(action_masks, n, device):
    """Crop to the logits width (generic/agent.py:202-203) and view the bools as bytes."""
    if action_masks is None:
        return None
    mk = torch.as_tensor(action_masks, device=device)
    if mk.dtype != torch.bool:
        mk = mk.to(torch.bool)
    mk = mk[:, :n]
    if mk.stride(-1) != 1:
        mk = mk.contiguous()
    return mk.view(torch.uint8)


@eager_only
def masked_log_softmax(logits, action_masks=None, fill=-1e20):
    """(logp [B,n], entropy [B], argmax [B]) of the masked categorical."""
    return _MaskedHead.apply(logits, _mask_u8(action_masks, logits.shape[1], logits.device), float(fill))


@eager_only
def masked_logprob_entropy(logits, action, action_masks=None, fill=-1e20):
    return _MaskedHeadEval.apply(logits, _mask_u8(action_masks, logits.shape[1], logits.device), action,
                                 float(fill))
