"""Dense projections of the path (``y = x @ W.T + b``): the node/edge embedding MLPs around every
conv (generic/mlp.py:70-77) and the attention-layer projections (generic/eegatconv.py:260-262,288),
SURVEY K1/K9, north_star item 4.

When TF32 matmuls are allowed - the reference's setting, ``torch.set_float32_matmul_precision("high")``
(psp/train_psp.py:47), i.e. ``torch.backends.cuda.matmul.allow_tf32`` - forward, input-gradient and
weight-gradient GEMMs run on the library's hand-written tcgen05 kernel (``b200gat_gemm_tf32``: TMA
operand staging, fp32 accumulators in TMEM, bias / GELU / GELU' / residual fused in the epilogue; the
three EGAT projections are one launch).  With TF32 off (fp32-exact matmuls, the parity tests' default)
or for shapes the kernel does not cover (width-1 head outputs, the 30-wide embedder input) the GEMM is
the plain library one (cuBLAS through torch).  ``WHEATLEY_B200_GEMM=cublas`` forces the library GEMM."""
import os

import torch

from . import _lib
from ._lib import eager_only


def gemm_backend() -> str:
    return os.environ.get("WHEATLEY_B200_GEMM", "tcgen05")


def tf32_path_enabled() -> bool:
    """The hand-written tcgen05 GEMM is used unless ``WHEATLEY_B200_GEMM=cublas``: with TF32 operands when
    ``torch.backends.cuda.matmul.allow_tf32`` is set, with 3xTF32 split arithmetic (fp32 matmul accuracy) when it is
    not (``_lib.sync_gemm_precision``)."""
    return gemm_backend() != "cublas"


def _ok(t):
    return t.is_cuda and t.dtype == torch.float32


def _desc(a_major, b_major, M, N, K, A, lda, B, ldb, D, ldd, *, a_parts=1, a_part_stride=0, d_parts=1,
          d_part_stride=0, bias=None, residual=None, epi=_lib.EPI_NONE, D2=None, Zin=None, ld2=0, a_rowsum=None):
    g = _lib.Gemm()
    g.a_major, g.b_major, g.M, g.N, g.K = a_major, b_major, M, N, K
    g.A, g.lda, g.a_parts, g.a_part_stride = _lib.ptr(A), lda, a_parts, a_part_stride
    g.B, g.ldb = _lib.ptr(B), ldb
    g.D, g.ldd, g.d_parts, g.d_part_stride = _lib.ptr(D), ldd, d_parts, d_part_stride
    g.bias, g.residual, g.ldr = _lib.ptr(bias), _lib.ptr(residual), (residual.stride(0) if residual is not None else 0)
    g.epilogue, g.D2, g.Zin, g.ld2 = epi, _lib.ptr(D2), _lib.ptr(Zin), ld2
    g.workspace, g.workspace_elems, g.a_rowsum = None, 0, _lib.ptr(a_rowsum)
    return g


def _run(g, device, what):
    ws_elems = int(_lib.lib.b200gat_gemm_workspace_elems(g))
    ws = None
    if ws_elems:
        ws = torch.empty(ws_elems, dtype=torch.float32, device=device)
        g.workspace, g.workspace_elems = ws.data_ptr(), ws_elems
    with _lib.device_guard(device):
        _lib.check(_lib.lib.b200gat_gemm_tf32(g, _lib.stream_ptr()), what)


def _fwd_supported(x, W):
    return (_ok(x) and _ok(W) and x.dim() == 2 and x.shape[0] > 0 and x.shape[1] % 4 == 0
            and W.shape[0] % 4 == 0)


@eager_only
def linear_fwd(x, W, b=None, gelu=False, residual=None, want_z=True):
    """``z = x W^T + b`` ; returns ``(z or None, out)`` with ``out = gelu(z)`` when ``gelu`` else
    ``z (+ residual)``.  x [M, K] contiguous, W [N, K]."""
    M, K = x.shape
    N = W.shape[0]
    if tf32_path_enabled() and _fwd_supported(x, W):
        W = W.contiguous()
        if gelu:
            z = torch.empty((M, N), dtype=torch.float32, device=x.device) if want_z else None
            h = torch.empty((M, N), dtype=torch.float32, device=x.device)
            g = _desc(0, 0, M, N, K, x, K, W, K, z, N, bias=b, epi=_lib.EPI_GELU, D2=h, ld2=N)
            _run(g, x.device, "b200gat_gemm_tf32 (linear+gelu)")
            return z, h
        y = torch.empty((M, N), dtype=torch.float32, device=x.device)
        g = _desc(0, 0, M, N, K, x, K, W, K, y, N, bias=b, residual=residual)
        _run(g, x.device, "b200gat_gemm_tf32 (linear)")
        return None, y
    z = torch.nn.functional.linear(x, W, b)
    if gelu:
        return z, torch.nn.functional.gelu(z)
    return None, (z if residual is None else z + residual)


@eager_only
def linear_dx(dy, W, z_prev=None):
    """``dx = dy W`` (``* gelu'(z_prev)`` when given).  dy [M, N], W [N, K] -> [M, K]."""
    M, N = dy.shape
    K = W.shape[1]
    if tf32_path_enabled() and _ok(dy) and _ok(W) and M > 0 and N % 4 == 0 and K % 4 == 0:
        W = W.contiguous()
        dx = torch.empty((M, K), dtype=torch.float32, device=dy.device)
        g = _desc(0, 1, M, K, N, dy, N, W, K, dx, K,
                  epi=_lib.EPI_DGELU if z_prev is not None else _lib.EPI_NONE, Zin=z_prev, ld2=K)
        _run(g, dy.device, "b200gat_gemm_tf32 (dx)")
        return dx
    dx = dy @ W
    if z_prev is not None:
        dx = torch.ops.aten.gelu_backward(dx, z_prev)
    return dx


@eager_only
def linear_dw(dy, x, want_bias=False):
    """``dW = dy^T x``: dy [M, N], x [M, K] -> [N, K] (reduction over the M rows, split-K); with
    ``want_bias`` also ``db = dy.sum(0)`` (row sums of the A operand, from the same launch).
    Returns ``(dW, db or None)``."""
    M, N = dy.shape
    K = x.shape[1]
    if tf32_path_enabled() and _ok(dy) and _ok(x) and M > 0 and N % 4 == 0 and K % 4 == 0:
        dW = torch.empty((N, K), dtype=torch.float32, device=dy.device)
        db = torch.empty(N, dtype=torch.float32, device=dy.device) if want_bias else None
        g = _desc(1, 1, N, K, M, dy, N, x, K, dW, K, a_rowsum=db)
        _run(g, dy.device, "b200gat_gemm_tf32 (dW)")
        return dW, db
    return dy.t() @ x, (colsum(dy) if want_bias else None)


@eager_only
def colsum(x):
    """Bias gradient ``x.sum(0)`` of a contiguous [M, N] tensor, deterministic."""
    M, N = x.shape
    if not (_ok(x) and tf32_path_enabled()):
        return x.sum(0)
    out = torch.empty(N, dtype=torch.float32, device=x.device)
    scratch = torch.empty(max(1, int(_lib.lib.b200gat_colsum_scratch_elems(M, N))), dtype=torch.float32,
                          device=x.device)
    with _lib.device_guard(x.device):
        _lib.check(_lib.lib.b200gat_colsum(_lib.ptr(x), M, N, x.stride(0), _lib.ptr(out), _lib.ptr(scratch),
                                           _lib.stream_ptr()), "b200gat_colsum")
    return out


class _GeluMLP(torch.autograd.Function):
    """``Linear -> GELU -> ... -> Linear (+ residual)`` with the activations fused into the GEMM
    epilogues (forward: bias + GELU; backward: GELU' applied to dx as it leaves TMEM)."""

    @staticmethod
    def forward(ctx, x, residual, *params):
        n = len(params) // 2
        need = any(ctx.needs_input_grad)
        saved, h = [], x
        for i in range(n - 1):
            z, h_next = linear_fwd(h, params[2 * i], params[2 * i + 1], gelu=True, want_z=need)
            saved += [h, z]
            h = h_next
        _, y = linear_fwd(h, params[2 * n - 2], params[2 * n - 1], residual=residual)
        saved.append(h)
        ctx.n = n
        ctx.has_res = residual is not None
        if need:
            ctx.save_for_backward(*saved, *params[0::2])
        return y

    @staticmethod
    def backward(ctx, gy):
        n = ctx.n
        saved = ctx.saved_tensors
        acts, Ws = saved[: 2 * n - 1], saved[2 * n - 1:]
        gy = gy.contiguous()
        grads = [None] * (2 * n)
        dz = gy
        for i in range(n - 1, -1, -1):
            h_in = acts[2 * i]                      # input of layer i
            if ctx.needs_input_grad[2 + 2 * i]:
                grads[2 * i], grads[2 * i + 1] = linear_dw(dz, h_in, want_bias=ctx.needs_input_grad[3 + 2 * i])
            elif ctx.needs_input_grad[3 + 2 * i]:
                grads[2 * i + 1] = colsum(dz)
            if i > 0:
                dz = linear_dx(dz, Ws[i], acts[2 * i - 1])      # * gelu'(z_{i-1})
            elif ctx.needs_input_grad[0]:
                dz = linear_dx(dz, Ws[0])
            else:
                dz = None
        return (dz, gy if ctx.has_res and ctx.needs_input_grad[1] else None, *grads)


def _mlp_desc(x, Ws, bs, residual, zs, hs, y):
    m = _lib.Mlp()
    n = len(Ws)
    m.n_layers, m.M = n, x.shape[0]
    m.dims[0] = x.shape[1]
    for i in range(n):
        m.dims[i + 1] = Ws[i].shape[0]
        m.W[i] = Ws[i].data_ptr()
        m.b[i] = None if bs[i] is None else bs[i].data_ptr()
    for i in range(n - 1):
        m.z[i] = None if zs is None else zs[i].data_ptr()
        m.h[i] = hs[i].data_ptr()
    m.x, m.residual, m.y = x.data_ptr(), _lib.ptr(residual), y.data_ptr()
    return m


def _pad4(n):
    return (n + 3) // 4 * 4


def _padded_rows_view(x):
    """``x [M, K]`` with ``K % 4 != 0`` whose rows already sit ``roundup4(K)`` floats apart in storage that holds the
    pad columns too (the trunk readout allocates its output that way, finite pad values): the ``[M, roundup4(K)]``
    matrix over the same storage - no copy - else None."""
    M, K = x.shape
    Kp = _pad4(K)
    if (x.stride(1) == 1 and x.stride(0) == Kp and x.data_ptr() % 16 == 0
            and (x.storage_offset() + (M - 1) * Kp + Kp) * 4 <= x.untyped_storage().nbytes()):
        return x.as_strided((M, Kp), (Kp, 1))
    return None


class _GeluMLPFused(torch.autograd.Function):
    """The whole ``Linear -> GELU -> ... -> Linear (+ residual)`` stack with ONE binding call per pass
    (``b200gat_mlp_fwd`` / ``b200gat_mlp_bwd``, which issue the tcgen05 GEMMs of ``_GeluMLP`` from C): the eager
    PPO step is host-bound and an MLP's forward + backward cost ~400 us of host time call by call.

    Widths that are not multiples of 4 floats (TMA rows must be 16-byte aligned: the 30-wide JSSP node features,
    the 478-wide head input, the width-1 head outputs) are zero-padded HERE, outside autograd (the padding used to
    be ``F.pad`` calls tracked by autograd: 20 launches + 10 backward nodes per step); the gradients are sliced back
    as views."""

    @staticmethod
    def forward(ctx, x, residual, *params):
        n = len(params) // 2
        Ws = [w.contiguous() for w in params[0::2]]
        bs = list(params[1::2])
        need = any(ctx.needs_input_grad)
        M, K, dev = x.shape[0], x.shape[1], x.device
        n_out = Ws[-1].shape[0]
        pad_in, pad_out = (-K) % 4, (-n_out) % 4
        if pad_in:
            xv = _padded_rows_view(x)
            x = xv if xv is not None else torch.nn.functional.pad(x, (0, pad_in))
            Ws[0] = torch.nn.functional.pad(Ws[0], (0, pad_in))
        elif not x.is_contiguous():
            x = x.contiguous()
        if pad_out:
            Ws[-1] = torch.nn.functional.pad(Ws[-1], (0, 0, 0, pad_out))
            if bs[-1] is not None:
                bs[-1] = torch.nn.functional.pad(bs[-1], (0, pad_out))
            if residual is not None:
                residual = torch.nn.functional.pad(residual, (0, pad_out))
        if residual is not None:
            residual = residual.contiguous()
        f32 = dict(dtype=torch.float32, device=dev)
        hs = [torch.empty((M, Ws[i].shape[0]), **f32) for i in range(n - 1)]
        zs = [torch.empty((M, Ws[i].shape[0]), **f32) for i in range(n - 1)] if need else None
        y = torch.empty((M, Ws[-1].shape[0]), **f32)
        m = _mlp_desc(x, Ws, bs, residual, zs, hs, y)
        with _lib.device_guard(dev):
            _lib.check(_lib.lib.b200gat_mlp_fwd(m, _lib.stream_ptr()), "b200gat_mlp_fwd")
        ctx.n, ctx.has_res = n, residual is not None
        ctx.has_bias = [b is not None for b in bs]
        ctx.K, ctx.n_out = K, n_out
        if need:
            ctx.save_for_backward(x, *Ws, *zs, *hs)
        return y[:, :n_out] if pad_out else y

    @staticmethod
    def backward(ctx, gy):
        n = ctx.n
        saved = ctx.saved_tensors
        x, Ws = saved[0], list(saved[1: 1 + n])
        rest = saved[1 + n:]
        zs, hs = list(rest[: n - 1]), list(rest[n - 1:])
        K, n_out = ctx.K, ctx.n_out
        pad_out = Ws[-1].shape[0] - n_out
        gy_in = gy
        if pad_out:
            gy = torch.nn.functional.pad(gy, (0, pad_out))
        gy = gy.contiguous()
        M, dev = x.shape[0], x.device
        f32 = dict(dtype=torch.float32, device=dev)
        need = ctx.needs_input_grad
        g = _lib.MlpGrads()
        g.gy = gy.data_ptr()
        dWs, dbs = [None] * n, [None] * n
        for i in range(n):
            if need[2 + 2 * i] or need[3 + 2 * i]:           # a bias gradient rides on its weight-gradient GEMM
                dWs[i] = torch.empty_like(Ws[i])
                g.dW[i] = dWs[i].data_ptr()
                if need[3 + 2 * i] and ctx.has_bias[i]:
                    dbs[i] = torch.empty(Ws[i].shape[0], **f32)
                    g.db[i] = dbs[i].data_ptr()
        dzs = [torch.empty_like(z) for z in zs]
        for i in range(n - 1):
            g.dz[i] = dzs[i].data_ptr()
        dx = torch.empty_like(x) if need[0] else None
        g.dx = _lib.ptr(dx)
        m = _mlp_desc(x, Ws, [None] * n, None, zs, hs, gy)       # biases / residual / y are not read backward
        ws_elems = int(_lib.lib.b200gat_mlp_bwd_workspace_elems(m, g))
        ws = torch.empty(max(ws_elems, 4), **f32)
        g.workspace, g.workspace_elems = ws.data_ptr(), ws_elems
        with _lib.device_guard(dev):
            _lib.check(_lib.lib.b200gat_mlp_bwd(m, g, _lib.stream_ptr()), "b200gat_mlp_bwd")
        # slice the padding off again (views)
        if dx is not None and dx.shape[1] != K:
            dx = dx[:, :K]
        if dWs[0] is not None and dWs[0].shape[1] != K:
            dWs[0] = dWs[0][:, :K]
        if pad_out:
            if dWs[-1] is not None:
                dWs[-1] = dWs[-1][:n_out]
            if dbs[-1] is not None:
                dbs[-1] = dbs[-1][:n_out]
        grads = []
        for i in range(n):
            grads += [dWs[i] if need[2 + 2 * i] else None, dbs[i] if need[3 + 2 * i] else None]
        return (dx, gy_in if ctx.has_res and need[1] else None, *grads)


@eager_only
def gelu_mlp(x, layers, residual=None):
    """``layers``: the ``nn.Linear`` modules of a GELU MLP.  One fused call per pass when every hidden width is a
    multiple of 4 floats (input / output widths are padded inside the call); GEMM by GEMM otherwise."""
    params = []
    for lin in layers:
        params += [lin.weight, lin.bias]
    hidden_ok = all(p.shape[0] % 4 == 0 for p in params[0:-2:2])
    if (hidden_ok and len(layers) <= _lib.MLP_MAX_LAYERS
            and os.environ.get("WHEATLEY_B200_MLP_CALLS", "fused") == "fused"):
        return _GeluMLPFused.apply(x, residual, *params)
    pad_in = (-x.shape[1]) % 4
    if pad_in:
        x = torch.nn.functional.pad(x, (0, pad_in))
        params[0] = torch.nn.functional.pad(params[0], (0, pad_in))
    n_out = params[-2].shape[0]
    pad_out = (-n_out) % 4
    if pad_out:
        params[-2] = torch.nn.functional.pad(params[-2], (0, 0, 0, pad_out))
        if params[-1] is not None:
            params[-1] = torch.nn.functional.pad(params[-1], (0, pad_out))
        if residual is not None:
            residual = torch.nn.functional.pad(residual, (0, pad_out))
    y = _GeluMLP.apply(x.contiguous(), None if residual is None else residual.contiguous(), *params)
    return y[:, :n_out] if pad_out else y


def gelu_mlp_applicable(x, layers) -> bool:
    return (tf32_path_enabled() and x.is_cuda and x.dtype == torch.float32 and x.dim() == 2 and x.shape[0] > 0)


class _Proj3(torch.autograd.Function):
    """The three bias-free projections of one attention layer (fc_ni | fc_nj | fc_node_src,
    generic/eegatconv.py:260-261,288) as ONE GEMM writing ``[3, M, HD]``; the backward reads the three
    incoming gradients as one k-concatenated operand (no copies when the conv backward wrote them into
    one buffer, which it does)."""

    @staticmethod
    def forward(ctx, x, w_i, w_j, w_v):
        M, K = x.shape
        HD = w_i.shape[0]
        wcat = torch.cat((w_i, w_j, w_v), 0)                       # [3 HD, K]
        p3 = torch.empty((3, M, HD), dtype=torch.float32, device=x.device)
        g = _desc(0, 0, M, 3 * HD, K, x, K, wcat, K, p3, HD, d_parts=3, d_part_stride=M * HD)
        _run(g, x.device, "b200gat_gemm_tf32 (projections)")
        ctx.save_for_backward(x, wcat)
        return p3[0], p3[1], p3[2]

    @staticmethod
    def backward(ctx, g_i, g_j, g_v):
        x, wcat = ctx.saved_tensors
        M, K = x.shape
        HD = wcat.shape[0] // 3
        gs = [g_i, g_j, g_v]
        if any(g is None for g in gs):
            gs = [torch.zeros((M, HD), dtype=torch.float32, device=x.device) if g is None else g for g in gs]
        step = M * HD * 4
        if (all(g.is_contiguous() for g in gs) and gs[1].data_ptr() == gs[0].data_ptr() + step
                and gs[2].data_ptr() == gs[1].data_ptr() + step):
            g3 = gs[0]                                             # base pointer of the shared [3, M, HD] buffer
        else:
            g3 = torch.stack(gs, 0)
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty((M, K), dtype=torch.float32, device=x.device)
            g = _desc(0, 1, M, K, 3 * HD, g3, HD, wcat, K, dx, K, a_parts=3, a_part_stride=M * HD)
            _run(g, x.device, "b200gat_gemm_tf32 (projections dx)")
        dw = torch.empty((3 * HD, K), dtype=torch.float32, device=x.device)
        g = _desc(1, 1, 3 * HD, K, M, g3, HD, x, K, dw, K, a_parts=3, a_part_stride=M * HD)
        _run(g, x.device, "b200gat_gemm_tf32 (projections dW)")
        return dx, dw[:HD], dw[HD: 2 * HD], dw[2 * HD:]


@eager_only
def project3(x, w_i, w_j, w_v):
    """``(x w_i^T, x w_j^T, x w_v^T)``; one tcgen05 launch when covered, three library GEMMs otherwise."""
    HD, K = w_i.shape
    if (tf32_path_enabled() and _ok(x) and x.dim() == 2 and x.shape[0] > 0 and K % 64 == 0 and HD % 128 == 0
            and w_j.shape == w_i.shape and w_v.shape == w_i.shape):
        return _Proj3.apply(x.contiguous(), w_i, w_j, w_v)
    lin = torch.nn.functional.linear
    return lin(x, w_i), lin(x, w_j), lin(x, w_v)


@eager_only
def dense_linear(x, weight, bias=None):
    """Single ``Linear`` (``lin_l`` / ``lin_r`` of the PyG-flavour layers, psp/graph/graph_conv.py:33-40):
    the one-layer case of the fused path (forward, dx and dW + bias row sums on ``b200gat_gemm_tf32``) when
    TF32 matmuls are allowed and the shape is covered, the library GEMM otherwise."""
    if (tf32_path_enabled() and _ok(x) and _ok(weight) and x.dim() == 2 and x.shape[0] > 0
            and x.shape[1] % 4 == 0 and weight.shape[0] % 4 == 0):
        return _GeluMLP.apply(x.contiguous(), None, weight, bias)
    return torch.nn.functional.linear(x, weight, bias)
