"""One PPO minibatch evaluated behind ONE autograd node: ``agent.get_action_and_value`` (generic/agent.py:190-231 -
trunk, ``value_net``, ``action_net``, masked categorical) + the clipped-PPO loss of alg/ppo.py:582-705, forward and
backward, one C-ABI call per pass (``b200gat_agent_fwd`` / ``b200gat_agent_bwd``, csrc/agent.cu).

This replaces the block alg/ppo.py:575-711 of the reference's update loop:

    _, newlogprob, entropy, newvalue, _ = agent.get_action_and_value(obs, action=..., action_masks=...)   # :575-581
    ... ratio / clipping / value loss / entropy, eight .item() read-backs ...                             # :582-710
    loss.backward()                                                                                       # :711

by

    loss, stats = agent.ppo_minibatch(obs, actions, masks, old_logprobs, advantages, returns, old_values, **cfg)
    loss.backward()

The module-level API (``get_action_and_value`` -> ``wheatley_b200.ppo.ppo_loss``) computes the same thing through ~10
autograd nodes; at 8 x Taillard 100x20 that path spends 5.6 ms of host time per step launching 4.5 ms of device work.
Same kernels, same results."""
import ctypes as C

import torch

from . import _lib
from ._lib import eager_only
from .readout import POOLING
from .trunk import _c, setup_trunk, setup_trunk_grads

STATS = ("pg_loss", "v_loss", "entropy_loss", "approx_kl", "clipfrac", "espo_dev", "adv_mean", "adv_std")


def _pad4(n):
    return (n + 3) // 4 * 4


class StepSpec:
    """Everything non-differentiable ``_AgentStepFn`` needs (built by the agent classes)."""

    def __init__(self, *, view, etype_sorted, attr_sorted, trunk_cfg, n_embedder, n_action, n_value, src_kinds,
                 src_widths, counts, row_start, max_n_nodes, pooling, pool_row0, ppo):
        self.view, self.etype_sorted, self.attr_sorted, self.trunk_cfg = view, etype_sorted, attr_sorted, trunk_cfg
        self.n_embedder, self.n_action, self.n_value = n_embedder, n_action, n_value
        self.src_kinds, self.src_widths = list(src_kinds), list(src_widths)
        self.counts, self.row_start, self.max_n_nodes = list(counts), row_start, int(max_n_nodes)
        self.pooling, self.pool_row0, self.ppo = pooling, int(pool_row0), dict(ppo)


def _mlp_struct(m, M, x_ptr, Ws, bs, y, zs, hs, residual=None):
    n = len(Ws)
    m.n_layers, m.M = n, M
    m.dims[0] = Ws[0].shape[1]
    for i in range(n):
        m.dims[i + 1] = Ws[i].shape[0]
        m.W[i] = Ws[i].data_ptr()
        m.b[i] = None if bs[i] is None else bs[i].data_ptr()
    for i in range(n - 1):
        m.h[i] = hs[i].data_ptr()
        m.z[i] = zs[i].data_ptr() if zs else None
    m.x, m.residual, m.y = x_ptr, _lib.ptr(residual), y.data_ptr()


class _AgentStepFn(torch.autograd.Function):
    """inputs: spec, ext (the external readout source = raw node features, or None), xin (embedder input, or the
    trunk input x0 when the agent embeds the nodes itself), action, mask_u8, old_logprob, advantages, returns,
    old_values, then the flat parameters: embedder (W, b)*, trunk layers (LayerSpec.flat order), action_net (W, b)*,
    value_net (W, b)*."""

    @staticmethod
    def forward(ctx, spec, ext, xin, action, mask_u8, old_lp, adv, ret, old_v, *flat):
        if not xin.is_cuda:
            raise _lib.B200GatError("wheatley_b200 kernels need CUDA tensors (no CPU fallback)")
        dev = xin.device
        f32 = dict(dtype=torch.float32, device=dev)
        tc = spec.trunk_cfg
        ne, na, nv = spec.n_embedder, spec.n_action, spec.n_value
        per = 2 * tc["n_proj"] + 4 + 2 * tc["n_mlp"]
        nt = tc["n_layers"] * per
        p_emb, p_trunk = list(flat[:2 * ne]), list(flat[2 * ne: 2 * ne + nt])
        p_act = list(flat[2 * ne + nt: 2 * ne + nt + 2 * na])
        p_val = list(flat[2 * ne + nt + 2 * na:])
        assert len(p_val) == 2 * nv
        need = any(ctx.needs_input_grad)
        a = _lib.AgentStep()
        keep, pads = [], []

        def pad_job(src, rows, cols, ld_dst, rows_dst=None):
            """16-byte aligned, zero-padded copy of a ragged 2-D operand, made by the call's first kernel."""
            rows_dst = rows if rows_dst is None else rows_dst
            dst = torch.empty((rows_dst, ld_dst), **f32)
            pads.append((src, dst, src.stride(0), ld_dst, rows, cols, rows_dst))
            return dst

        def padded_params(params, k_in):
            """(Ws, bs) of an MLP with the first weight's columns padded to ``k_in`` and the last layer's rows (and
            bias) padded to a multiple of 4."""
            Ws = [_c(w) for w in params[0::2]]
            bs = [None if b is None else _c(b) for b in params[1::2]]
            if Ws[0].shape[1] != k_in:
                Ws[0] = pad_job(Ws[0], Ws[0].shape[0], Ws[0].shape[1], k_in)
            n_out = Ws[-1].shape[0]
            if n_out % 4:                       # width-1 head outputs: rows (and bias entries) padded to 4
                rows_p = _pad4(n_out)
                Ws[-1] = pad_job(Ws[-1], n_out, Ws[-1].shape[1], Ws[-1].shape[1], rows_p)
                if bs[-1] is not None:
                    bs[-1] = pad_job(bs[-1].reshape(1, -1), 1, n_out, rows_p).reshape(-1)
            return Ws, bs

        # ---- embedder (jssp/models/gnn_dgl.py:355) or a caller-built x0
        N = xin.shape[0]
        if ne:
            F_in = xin.shape[1]
            K_in = _pad4(F_in)
            x_src = _c(xin)
            x_emb = pad_job(x_src, N, F_in, K_in) if K_in != F_in else x_src
            We, be = padded_params(p_emb, K_in)
            hidden = We[-1].shape[0]
            x0 = torch.empty((N, hidden), **f32)
            he = [torch.empty((N, w.shape[0]), **f32) for w in We[:-1]]
            ze = [torch.empty((N, w.shape[0]), **f32) for w in We[:-1]] if need else []
            a.has_embedder = 1
            _mlp_struct(a.embedder, N, x_emb.data_ptr(), We, be, x0, ze, he)
            keep += [x_src, x_emb, *We, *be, *he, *ze]
        else:
            x0 = _c(xin)
            a.has_embedder = 0
        # ---- trunk
        flat_c, tkeep, y = setup_trunk(a.trunk, spec.view, tc, spec.etype_sorted, spec.attr_sorted, x0, p_trunk, need)
        keep += tkeep
        # ---- readout
        r = a.readout
        B, max_n = len(spec.counts), spec.max_n_nodes
        r.n_src = len(spec.src_kinds)
        ext_c = None
        for i, (kind, w) in enumerate(zip(spec.src_kinds, spec.src_widths)):
            if kind == _lib.SRC_EXTERNAL:
                ext_c = ext if ext.stride(1) == 1 else ext.contiguous()
                r.src[i], r.ld[i] = ext_c.data_ptr(), ext_c.stride(0)
            elif kind == _lib.SRC_X0:
                r.src[i], r.ld[i] = x0.data_ptr(), x0.stride(0)
            else:
                r.src[i], r.ld[i] = y.data_ptr() + (kind - _lib.SRC_LAYER0) * y.stride(0) * 4, y.stride(1)
            r.width[i] = w
            a.src_kind[i] = kind
        F = sum(spec.src_widths)
        Fp = _pad4(F)
        mode = POOLING[spec.pooling]
        r.n_graphs, r.max_n_nodes, r.row_start = B, max_n, spec.row_start.data_ptr()
        r.n_task_rows, r.pooling, r.pool_row0, r.out_ld = sum(spec.counts), mode, spec.pool_row0, Fp
        nf = torch.empty((B, max_n, Fp), **f32)
        ge = torch.empty((B, Fp), **f32)
        am = torch.empty((B, F), dtype=torch.int32, device=dev) if mode == _lib.READOUT_MAX else None
        a.node_features, a.graph_embedding, a.argmax = nf.data_ptr(), ge.data_ptr(), _lib.ptr(am)
        # ---- heads (generic/agent.py:194-195)
        Ma = B * max_n
        Wa, ba = padded_params(p_act, Fp)
        Wv, bv = padded_params(p_val, Fp)
        ya = torch.empty((Ma, Wa[-1].shape[0]), **f32)
        yv = torch.empty((B, Wv[-1].shape[0]), **f32)
        ha = [torch.empty((Ma, w.shape[0]), **f32) for w in Wa[:-1]]
        za = [torch.empty((Ma, w.shape[0]), **f32) for w in Wa[:-1]] if need else []
        hv = [torch.empty((B, w.shape[0]), **f32) for w in Wv[:-1]]
        zv = [torch.empty((B, w.shape[0]), **f32) for w in Wv[:-1]] if need else []
        _mlp_struct(a.action_net, Ma, nf.data_ptr(), Wa, ba, ya, za, ha)
        _mlp_struct(a.value_net, B, ge.data_ptr(), Wv, bv, yv, zv, hv)
        # ---- masked head + loss
        action = action.long().contiguous()
        f = lambda t_: None if t_ is None else t_.detach().to(torch.float32).reshape(-1).contiguous()   # noqa: E731
        old_lp, adv, ret, old_v = f(old_lp), f(adv), f(ret), f(old_v)
        small = torch.empty(2 * Ma + 6 * B + 12, **f32)          # logits | logp | head [3,B] | g_head [3,B] | loss [9]
        logits, logp = small[:Ma], small[Ma: 2 * Ma]
        head = small[2 * Ma: 2 * Ma + 3 * B]
        g_head = small[2 * Ma + 3 * B: 2 * Ma + 6 * B]
        loss_out = small[2 * Ma + 6 * B: 2 * Ma + 6 * B + 9]
        a.mask, a.ld_mask = _lib.ptr(mask_u8), (0 if mask_u8 is None else mask_u8.stride(0))
        a.action, a.fill = action.data_ptr(), -1e20
        a.logits, a.logp, a.head = logits.data_ptr(), logp.data_ptr(), head.data_ptr()
        a.loss_out, a.g_head = loss_out.data_ptr(), g_head.data_ptr()
        ppo = spec.ppo
        a.loss = _lib.PpoLossArgs(B, None, _lib.ptr(old_lp), _lib.ptr(adv), None, None, _lib.ptr(ret),
                                  _lib.ptr(old_v), -1.0 if ppo["clip_coef"] is None else ppo["clip_coef"],
                                  int(ppo["norm_adv"]), int(ppo["clip_vloss"]), _lib.CRITIC[ppo["critic_loss"]],
                                  ppo["ent_coef"], ppo["vf_coef"], _lib.ptr(ppo.get("moments")))
        # ---- the pad jobs
        if len(pads) > _lib.AGENT_MAX_PAD:
            raise _lib.B200GatError("too many ragged operands for one agent step")
        for k, (src, dst, ld_src, ld_dst, rows, cols, rows_dst) in enumerate(pads):
            a.pad[k] = _lib.PadJob(src.data_ptr(), dst.data_ptr(), ld_src, ld_dst, rows, cols, rows_dst, 0)
            keep.append(src)
        a.n_pad = len(pads)
        with _lib.device_guard(dev):
            _lib.check(_lib.lib.b200gat_agent_fwd(C.byref(a), _lib.stream_ptr()), "b200gat_agent_fwd")
        if need:
            ctx.a, ctx.spec = a, spec
            ctx.sizes = (ne, nt, na, nv, N, F, Fp)
            ctx.shapes = [None if p is None else tuple(p.shape) for p in flat]
            ctx.n_flat_trunk = len(flat_c)
            ctx.has_ext = ext is not None
            ctx.ext_shape = None if ext is None else tuple(ext.shape)
            ctx.xin_shape = tuple(xin.shape)
            # every tensor the struct points into stays alive with the node
            ctx.save_for_backward(*[t_ for t_ in keep + [ext_c, x0, nf, ge, am, ya, yv, *ha, *za, *hv, *zv, *Wa,
                                                          *[b_ for b_ in ba if b_ is not None], *Wv,
                                                          *[b_ for b_ in bv if b_ is not None], action, mask_u8,
                                                          old_lp, adv, ret, old_v, small, spec.row_start,
                                                          ppo.get("moments")]
                                    if isinstance(t_, torch.Tensor)])
            ctx.flat_c = None
            ctx.trunk_flat = flat_c
        stats = loss_out[1:]
        aux = head.view(3, B)
        ctx.mark_non_differentiable(stats, aux)
        return loss_out[0], stats, aux

    @staticmethod
    def backward(ctx, g_loss, _g_stats, _g_aux):
        a, spec = ctx.a, ctx.spec
        ne, nt, na, nv, N, F, Fp = ctx.sizes
        tc = spec.trunk_cfg
        n_in = 9                                     # non-parameter inputs of forward
        n_flat = 2 * ne + nt + 2 * na + 2 * nv
        if g_loss is None:
            return (None,) * (n_in + n_flat)
        dev = g_loss.device
        f32 = dict(dtype=torch.float32, device=dev)
        g = _lib.AgentStepGrads()
        g_loss = g_loss.detach().to(torch.float32).reshape(1).contiguous()
        g.g_loss = g_loss.data_ptr()
        grads = [None] * n_flat

        def mlp_grads(mg, m, base, n, shapes):
            """Padded-shape gradient buffers for an MLP whose struct ``m`` holds the padded widths; the returned
            gradients are views with the parameters' own shapes."""
            for i in range(n):
                rows, cols = m.dims[i + 1], m.dims[i]
                dW = torch.empty((rows, cols), **f32)
                mg.dW[i] = dW.data_ptr()
                ws_, bs_ = shapes[base + 2 * i], shapes[base + 2 * i + 1]
                grads[base + 2 * i] = dW[: ws_[0], : ws_[1]] if (rows, cols) != tuple(ws_) else dW
                if bs_ is not None:
                    db = torch.empty(rows, **f32)
                    mg.db[i] = db.data_ptr()
                    grads[base + 2 * i + 1] = db[: bs_[0]] if rows != bs_[0] else db

        shapes = ctx.shapes
        g_xin_emb = None
        if ne:
            mlp_grads(g.embedder, a.embedder, 0, ne, shapes)
            if ctx.needs_input_grad[2]:
                g_xin_emb = torch.empty((N, a.embedder.dims[0]), **f32)
                g.embedder.dx = g_xin_emb.data_ptr()
        tgrads = setup_trunk_grads(g.trunk, tc, ctx.trunk_flat, shapes[2 * ne: 2 * ne + nt], dev)
        grads[2 * ne: 2 * ne + nt] = tgrads
        mlp_grads(g.action_net, a.action_net, 2 * ne + nt, na, shapes)
        mlp_grads(g.value_net, a.value_net, 2 * ne + nt + 2 * na, nv, shapes)
        g_ext = None
        if ctx.has_ext and ctx.needs_input_grad[1]:
            g_ext = torch.empty(ctx.ext_shape, **f32)
            for i, kind in enumerate(spec.src_kinds):
                if kind == _lib.SRC_EXTERNAL:
                    g.g_src_ext[i] = g_ext.data_ptr()
        g_x0 = None
        if not ne and ctx.needs_input_grad[2]:
            g_x0 = torch.empty(ctx.xin_shape, **f32)
            g.g_x0 = g_x0.data_ptr()
        ws_elems = int(_lib.lib.b200gat_agent_bwd_workspace_elems(C.byref(a)))
        if ws_elems <= 0:
            raise _lib.B200GatError(f"b200gat_agent_bwd_workspace_elems: {_lib.lib.b200gat_last_error().decode()}")
        ws = torch.empty(ws_elems, **f32)
        g.workspace, g.workspace_elems = ws.data_ptr(), ws_elems
        with _lib.device_guard(dev):
            _lib.check(_lib.lib.b200gat_agent_bwd(C.byref(a), C.byref(g), _lib.stream_ptr()), "b200gat_agent_bwd")
        need = ctx.needs_input_grad[n_in:]
        grads = [gr if nd else None for gr, nd in zip(grads, need)]
        g_xin = g_x0
        if ne and g_xin_emb is not None:
            g_xin = g_xin_emb[:, : ctx.xin_shape[1]] if g_xin_emb.shape[1] != ctx.xin_shape[1] else g_xin_emb
        return (None, g_ext, g_xin, None, None, None, None, None, None, *grads)


@eager_only
def agent_step(spec, ext, xin, action, mask_u8, old_logprob, advantages, returns, old_values, flat):
    """``(loss, stats [8], aux [3, B] = logprob | entropy | value)``; ``stats`` in the order of ``STATS``."""
    return _AgentStepFn.apply(spec, ext, xin, action, mask_u8, old_logprob, advantages, returns, old_values, *flat)
