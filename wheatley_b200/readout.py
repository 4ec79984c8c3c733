"""Trunk readout: concat-all-layers + per-graph pooling + pad-to-``max_n_nodes`` as ONE pass over the per-layer
node tensors (``b200gat_readout_fwd/bwd``, csrc/readout.cu).

Replaces ``torch.cat(features_list)`` + the per-graph Python loops of psp/models/gnn_mp.py:419-500 and the
``cat`` / reshape / pool / ``F.pad`` tail of jssp/models/gnn_dgl.py:401-425."""
import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import eager_only

POOLING = {None: _lib.READOUT_NONE, "none": _lib.READOUT_NONE, "max": _lib.READOUT_MAX, "avg": _lib.READOUT_AVG,
           "learn": _lib.READOUT_ROWS, "learninv": _lib.READOUT_ROWS}


def row_starts(counts, device):
    """``[B+1]`` int32 device tensor: exclusive prefix of the per-graph task-node counts (host knowledge: the
    observation carries ``n_nodes``), one small copy per batch."""
    counts = np.asarray(counts, dtype=np.int64)
    if counts.size and (counts == counts[0]).all():          # equal-sized graphs: generated on the device
        n, B = int(counts[0]), len(counts)
        if n == 0:
            return torch.zeros(B + 1, dtype=torch.int32, device=device)
        return torch.arange(0, (B + 1) * n, n, dtype=torch.int32, device=device)
    rs = torch.zeros(len(counts) + 1, dtype=torch.int32, pin_memory=True)      # pinned: legal under graph capture
    np.cumsum(counts, out=rs.numpy()[1:])
    return rs.to(device, non_blocking=True)


class _Readout(torch.autograd.Function):
    @staticmethod
    def forward(ctx, row_start, B, max_n_nodes, n_task_rows, mode, pool_row0, *sources):
        if len(sources) > _lib.READOUT_MAX_SRC:
            raise _lib.B200GatError(f"at most {_lib.READOUT_MAX_SRC} sources")
        src = []
        for s in sources:
            if not s.is_cuda or s.dtype != torch.float32 or s.dim() != 2:
                raise _lib.B200GatError("readout sources must be 2-D fp32 CUDA tensors (no CPU fallback)")
            src.append(s if s.stride(1) == 1 else s.contiguous())
        n_rows = src[0].shape[0]
        if any(s.shape[0] != n_rows for s in src):
            raise _lib.B200GatError("readout sources must have the same number of rows")
        widths = [int(s.shape[1]) for s in src]
        F = sum(widths)
        dev = src[0].device
        r = _lib.Readout()
        r.n_src = len(src)
        for i, s in enumerate(src):
            r.src[i], r.width[i], r.ld[i] = s.data_ptr(), widths[i], s.stride(0)
        r.n_graphs, r.max_n_nodes, r.row_start = B, max_n_nodes, row_start.data_ptr()
        r.n_task_rows, r.pooling, r.pool_row0 = n_task_rows, mode, pool_row0
        # rows 16-byte aligned (zero pad columns; the returned tensors are the [..., :F] views): the heads' first
        # GEMM reads them in place instead of through a padded 30 MB copy
        Fp = (F + 3) // 4 * 4
        r.out_ld = Fp
        nf = torch.empty((B, max_n_nodes, Fp), dtype=torch.float32, device=dev)
        ge = torch.empty((B, Fp), dtype=torch.float32, device=dev) if mode != _lib.READOUT_NONE else None
        am = torch.empty((B, F), dtype=torch.int32, device=dev) if mode == _lib.READOUT_MAX else None
        with _lib.device_guard(dev):
            _lib.check(_lib.lib.b200gat_readout_fwd(C.byref(r), _lib.ptr(nf), _lib.ptr(ge), _lib.ptr(am),
                                                    _lib.stream_ptr()), "b200gat_readout_fwd")
        ctx.set_materialize_grads(False)
        ctx.meta = (B, max_n_nodes, n_task_rows, mode, pool_row0, widths, n_rows)
        ctx.save_for_backward(row_start, am)
        if ge is None:
            ge = nf.new_zeros((B, 0))
            ctx.mark_non_differentiable(ge)
        elif Fp != F:
            ge = ge[:, :F]
        return (nf[:, :, :F] if Fp != F else nf), ge

    @staticmethod
    def backward(ctx, g_nf, g_ge):
        B, max_n_nodes, n_task_rows, mode, pool_row0, widths, n_rows = ctx.meta
        row_start, am = ctx.saved_tensors
        need = ctx.needs_input_grad[6:]
        if g_nf is None and g_ge is None:
            return (None,) * (6 + len(widths))
        dev = row_start.device
        F = sum(widths)
        ld_nf = ld_ge = F
        if g_nf is not None:      # a gradient that arrives with padded rows (the head's dx) is read in place
            if g_nf.stride(2) == 1 and g_nf.stride(0) == max_n_nodes * g_nf.stride(1) and g_nf.stride(1) >= F:
                ld_nf = g_nf.stride(1)
            else:
                g_nf = g_nf.contiguous()
        if g_ge is not None:
            if g_ge.stride(1) == 1 and g_ge.stride(0) >= F:
                ld_ge = g_ge.stride(0)
            else:
                g_ge = g_ge.contiguous()
        grads = [torch.empty((n_rows, w), dtype=torch.float32, device=dev) if nd else None
                 for w, nd in zip(widths, need)]
        r = _lib.Readout()
        r.n_src = len(widths)
        for i, w in enumerate(widths):
            r.src[i], r.width[i], r.ld[i] = None, w, w
        r.n_graphs, r.max_n_nodes, r.row_start = B, max_n_nodes, row_start.data_ptr()
        r.n_task_rows, r.pooling, r.pool_row0 = n_task_rows, mode, pool_row0
        gp = (C.c_void_p * _lib.READOUT_MAX_SRC)(*[_lib.ptr(g) for g in grads])
        with _lib.device_guard(dev):
            _lib.check(_lib.lib.b200gat_readout_bwd(C.byref(r), _lib.ptr(g_nf), ld_nf, _lib.ptr(g_ge), ld_ge,
                                                    _lib.ptr(am), gp, n_rows, _lib.stream_ptr()),
                       "b200gat_readout_bwd")
        return (None,) * 6 + tuple(grads)


@eager_only
def readout(sources, counts, max_n_nodes, pooling, pool_row0=0, row_start=None):
    """``(node_features [B, max_n_nodes, F], graph_embedding [B, F] or None)`` from the per-layer node tensors
    ``sources`` (each ``[N_total, w_i]``; their concatenation along dim 1 is never materialised).

    ``counts``: task nodes per graph (host ints; the graphs' task rows are consecutive from row 0);
    ``pooling``: ``"max"`` / ``"avg"`` over each graph's task rows, ``"learn"`` / ``"learninv"``: row
    ``pool_row0 + b`` (the learned pool node of graph ``b``), ``None``: no graph embedding."""
    counts = [int(c) for c in counts]
    if counts and max(counts) > max_n_nodes:
        raise _lib.B200GatError(f"a graph has {max(counts)} nodes, max_n_nodes is {max_n_nodes}")
    mode = POOLING[pooling]
    dev = sources[0].device
    if row_start is None:
        row_start = row_starts(counts, dev)
    nf, ge = _Readout.apply(row_start, len(counts), int(max_n_nodes), int(sum(counts)), mode, int(pool_row0),
                            *sources)
    return nf, (None if mode == _lib.READOUT_NONE else ge)
