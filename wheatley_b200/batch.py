"""Device-side JSSP batch builder: the padded gym observation goes to the GPU as it is and the
batched DGL-flavour graph - same nodes, same edges IN THE SAME ORDER, same edge types and fixed-up features as
``AgentObservation.from_gym_observation`` (jssp/models/agent_observation.py:110-229) followed by the pool-node
wiring of ``GnnDGL.forward`` (jssp/models/gnn_dgl.py:263-286) - is written there by ``b200gat_jssp_batch_build``.

The reference builds every minibatch with Python loops over its graphs on the CPU ("batching on CPU for
performance reasons", agent_observation.py:126) and then copies the batched graph; at 8 x Taillard 100x20 that
graph is 41 MB of int64 COO (1.65 M edges, almost all of them machine-clique conflicts that are a function of
2,000 machine ids per graph), the observation it was built from 2.2 MB.

Covered: ``conflicts in ("att", "clique", "node")`` (``"node"``: machine nodes and their wiring,
gnn_dgl.py:288-335), ``graph_pooling in ("learn", "learninv", other = no pool node)``, ``bidir``,
``add_self_loops``, ``mid_in_edges``, ``no_tct``.  No CPU path: CUDA device required."""
import torch

from . import _lib
from .plan import GraphPlan


def jssp_sizes(features, n_nodes, n_edges, conflicts="clique", add_self_loops=True, bidir=True):
    """Host-side sizes of the batch from the observation alone (no device work): ``meta [4, B+1]`` int64
    (node offsets, edge offsets of the per-graph segments, precedence edges, clique edges per graph) and the
    number of distinct machine ids.  The clique edge count of a graph is ``sum_m c_m (c_m - 1)`` over its
    machine histogram (jssp/utils/utils.py:460-478 emits every ordered pair of one machine)."""
    import numpy as np
    B, n = int(features.shape[0]), int(features.shape[1])
    as_np = lambda x: (x.detach().cpu().numpy() if isinstance(x, torch.Tensor) else np.asarray(x))  # noqa: E731
    nn = as_np(n_nodes).astype(np.int64).reshape(-1)
    ne = as_np(n_edges).astype(np.int64).reshape(-1)
    # pinned: the copy to the device is asynchronous (a pageable source makes the copy synchronise its stream)
    meta = torch.zeros((4, B + 1), dtype=torch.int64, pin_memory=torch.cuda.is_available())
    m = meta.numpy()
    n_mid = 0
    K = np.zeros(B, np.int64)
    affected = np.zeros(B, np.int64)
    if conflicts in ("clique", "node") and B and n:
        mid = as_np(features)[:, :, 6].astype(np.int64)                          # .long(): truncation
        valid = (np.arange(n)[None, :] < nn[:, None]) & (mid >= 0)
        if valid.any():
            n_mid = int(mid[valid].max()) + 1
            flat = (np.arange(B)[:, None] * n_mid + mid)[valid]
            hist = np.bincount(flat, minlength=B * n_mid).reshape(B, n_mid)
            if conflicts == "clique":
                K = (hist * (hist - 1)).sum(1)
            else:       # "node": row 3 = exclusive prefix of the affected-task counts ([B] = total)
                affected = hist.sum(1)
    per_graph = (2 if bidir else 1) * ne + K + (nn if add_self_loops else 0)
    m[0, 1:] = np.cumsum(nn)
    m[1, 1:] = np.cumsum(per_graph)
    m[2, :B] = ne
    if conflicts == "node":
        m[3, 1:] = np.cumsum(affected) if n_mid else 0
    else:
        m[3, :B] = K
    return meta, n_mid


def jssp_batch_on_device(obs, device, conflicts="att", graph_pooling="learn", add_self_loops=True, bidir=True,
                         no_tct=False, mid_in_edges=False, max_n_machines=0):
    """``obs``: the gym observation of B graphs - ``features [B, max_nodes, F]`` fp32, ``edge_index
    [B, 2, max_edges]`` int64, ``n_nodes [B]``, ``n_edges [B]`` - as CPU tensors (pinned: the copies are
    asynchronous).  Returns ``dict(src, dst, etype [E] int64, feat [N (+B), F] fp32, n_nodes, n_task_nodes,
    poolnodes, batch_num_nodes)`` on ``device``, everything queued on the current stream."""
    device = torch.device(device)
    if device.type != "cuda":
        raise _lib.B200GatError("jssp_batch_on_device needs a CUDA device: wheatley_b200 has no CPU path")
    if conflicts not in ("att", "clique", "node"):
        raise ValueError(f"unknown conflicts mode {conflicts!r}")
    if conflicts == "node" and max_n_machines <= 0:
        raise ValueError("conflicts='node' needs max_n_machines (machine nodes per graph, gnn_dgl.py:288-335)")
    features, edge_index = obs["features"], obs["edge_index"]
    if features.dtype != torch.float32:
        features = features.float()
    if edge_index.dtype != torch.int64:
        edge_index = edge_index.long()
    features, edge_index = features.contiguous(), edge_index.contiguous()
    B, max_nodes, F = (int(x) for x in features.shape)
    max_edges = int(edge_index.shape[2]) if edge_index.dim() == 3 else 0
    meta, n_mid = jssp_sizes(features, obs["n_nodes"], obs["n_edges"], conflicts, add_self_loops, bidir)
    m = meta.numpy()
    if B and (int(m[2, :B].max()) > max_edges or int((m[0, 1:] - m[0, :-1]).max()) > max_nodes):
        raise _lib.B200GatError("n_nodes / n_edges exceed the padded observation arrays")
    pooling = _lib.POOLING.get(graph_pooling if graph_pooling in ("learn", "learninv") else None)
    N, Eg = int(m[0, B]), int(m[1, B])
    node_mode = conflicts == "node"
    M = int(max_n_machines) if node_mode else 0
    if node_mode and n_mid > M:
        raise _lib.B200GatError(f"machine id {n_mid - 1} with max_n_machines={M}")
    n_affected = int(m[3, B]) if node_mode else 0
    with _lib.device_guard(device):
        d_feat = features.to(device, non_blocking=True)
        d_ei = edge_index.to(device, non_blocking=True)
        d_meta = meta.to(device, non_blocking=True)
        o = _lib.JsspObs(B, max_nodes, F, max_edges, _lib.ptr(d_feat), _lib.ptr(d_ei), _lib.ptr(d_meta),
                         int(conflicts == "clique"), int(mid_in_edges), int(bidir), int(add_self_loops), int(no_tct),
                         pooling, n_mid, N, Eg, int(node_mode), M, n_affected)
        E = int(_lib.lib.b200gat_jssp_batch_edges(o))
        i64 = dict(dtype=torch.int64, device=device)
        src, dst, etype = torch.empty(E, **i64), torch.empty(E, **i64), torch.empty(E, **i64)
        n_total = N + (B if pooling else 0) + M * B
        feat = torch.empty((n_total, F), dtype=torch.float32, device=device)
        _lib.check(_lib.lib.b200gat_jssp_batch_build(o, _lib.ptr(src), _lib.ptr(dst), _lib.ptr(etype), _lib.ptr(feat),
                                                     _lib.stream_ptr()), "b200gat_jssp_batch_build")
        poolnodes = torch.arange(N, N + B, device=device) if pooling else None
        machinenodes = torch.arange(n_total - M * B, n_total, device=device) if node_mode else None
    return dict(src=src, dst=dst, etype=etype, feat=feat, n_nodes=n_total, n_task_nodes=N, poolnodes=poolnodes,
                machinenodes=machinenodes,
                batch_num_nodes=(m[0, 1:] - m[0, :-1]).tolist(), n_edges=E)


def jssp_batch_from_observation(obs, device, max_n_nodes, conflicts="att", graph_pooling="learn", defer=False, **kw):
    """Observation -> ``JsspBatch`` ready for ``GnnDGL.forward`` (the ``AgentObservation.from_gym_observation`` +
    ``to_graph`` step of jssp/models/gnn_dgl.py:239-262 and the wiring of :263-286), all on the device.
    ``defer=True`` leaves the plan's host counts pending (``JsspBatch.plan.finish()``, used by the prefetcher)."""
    from .gnn_dgl import JsspBatch
    d = jssp_batch_on_device(obs, device, conflicts, graph_pooling, **kw)
    bnn = d["batch_num_nodes"]
    # nodes per graph = num_nodes / batch_size (jssp/models/agent_observation.py:47-51); the trunk reshapes the task
    # rows to [B, n_nodes, F] (gnn_dgl.py:402-405: equal-sized graphs) and THEN pads to max_n_nodes (:420-425)
    if len(set(bnn)) > 1:
        raise _lib.B200GatError(f"graphs of one batch must have the same number of nodes (got {sorted(set(bnn))}): "
                                "GnnDGL reshapes the node rows to [B, n_nodes, F] (jssp/models/gnn_dgl.py:402-405)")
    n_nodes = int(bnn[0]) if bnn else 0
    if n_nodes > max_n_nodes:
        raise _lib.B200GatError(f"graphs have {n_nodes} nodes, max_n_nodes is {max_n_nodes}")
    plan = GraphPlan(d["src"], d["dst"], d["n_nodes"])
    plan.prepare(d["etype"], defer=defer)
    return JsspBatch(plan, d["etype"], d["feat"], len(bnn), n_nodes, d["n_task_nodes"],
                     poolnodes=d["poolnodes"], machinenodes=d["machinenodes"], prepare=not defer), d
