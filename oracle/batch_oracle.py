"""numpy restatement of the DGL-flavour batch construction.
TEST INFRASTRUCTURE - see oracle/__init__.py.  Pinned by tests/golden/jssp_batch.npz
(jssp/models/agent_observation.py executed verbatim on the DGL shim).

Integer work: bit-exact."""
import numpy as np


def conflicts_cliques(machineid: np.ndarray):
    """jssp/utils/utils.py:460-478: all ordered pairs (i,j), i != j, with the same
    machine id (unaffected tasks, id -1, conflict with nobody); row-major order of the
    (j,i) matrix ``cond[r,c] = (mid[c] == mid2[r]) & (r != c)``; returns
    ``(edges [2,K], machineid[edges[0]])``."""
    m = np.asarray(machineid)
    m1 = m[None, :]
    m2 = np.where(m == -1, -2, m)[:, None]
    cond = (m1 == m2) & ~np.eye(m.shape[0], dtype=bool)
    r, c = np.nonzero(cond)
    return np.stack([r, c]).astype(np.int64), m[r].astype(np.int64)


def jssp_batch(features, edge_index, n_nodes, n_edges, conflicts="att", mid_in_edges=False,
               add_self_loops=True, bidir=True, no_tct=False):
    """jssp/models/agent_observation.py:110-229.  Inputs are the padded gym observation
    arrays.  Returns dict(src, dst, type (int64 [E]), feat [N,F], batch_num_nodes,
    batch_num_edges)."""
    feat = np.array(features, dtype=np.float32, copy=True)
    feat[:, :, 1] = np.where(feat[:, :, 1] == -1, -1, feat[:, :, 4])     # :133-135
    if no_tct:
        feat[:, :, 1:5] = -1                                             # :138-139
    srcs, dsts, types, feats, bnn, bne = [], [], [], [], [], []
    off = 0
    for i in range(feat.shape[0]):
        nn, ne = int(n_nodes[i]), int(n_edges[i])
        e0, e1 = edge_index[i, 0, :ne], edge_index[i, 1, :ne]
        if bidir:                                                        # :57-66
            s, d = np.concatenate([e0, e1]), np.concatenate([e1, e0])
            t = np.concatenate([np.full(ne, 1), np.full(ne, 2)])
        else:
            s, d, t = e0, e1, np.full(ne, 1)
        if conflicts == "clique":                                        # :160-176, :81-88
            ce, cemid = conflicts_cliques(feat[i, :nn, 6].astype(np.int64))
            s, d = np.concatenate([s, ce[0]]), np.concatenate([d, ce[1]])
            t = np.concatenate([t, cemid + 5 if mid_in_edges else np.full(ce.shape[1], 5)])
        if add_self_loops:                                               # :210-211
            loop = np.arange(nn)
            s, d = np.concatenate([s, loop]), np.concatenate([d, loop])
            t = np.concatenate([t, np.zeros(nn, dtype=np.int64)])
        srcs.append(s + off)                                             # dgl.batch :223
        dsts.append(d + off)
        types.append(t)
        feats.append(feat[i, :nn])
        bnn.append(nn)
        bne.append(s.shape[0])
        off += nn
    return dict(src=np.concatenate(srcs).astype(np.int64), dst=np.concatenate(dsts).astype(np.int64),
                type=np.concatenate(types).astype(np.int64), feat=np.concatenate(feats),
                batch_num_nodes=np.array(bnn), batch_num_edges=np.array(bne))


def jssp_add_pool_and_machine_nodes(b, graph_pooling="learn", conflicts="att", max_n_machines=0):
    """jssp/models/gnn_dgl.py:263-335: pool nodes (one per graph) appended after all
    graph nodes; edges appended in order: pool self loops (type 0), [pool->node type 3
    if learninv], node->pool type 4; ``conflicts == 'node'``: machine nodes + self loops
    + task->machine (mid+5) + machine->task (mid+5+M) [+ machine->pool, type 5+2M]."""
    src, dst, typ, feat = b["src"], b["dst"], b["type"], b["feat"]
    bnn = b["batch_num_nodes"]
    B = len(bnn)
    num_nodes = feat.shape[0]
    node_offset = num_nodes
    out = dict(b)
    S, D, T = [src], [dst], [typ]
    if graph_pooling in ("learn", "learninv"):
        feat = np.concatenate([feat, np.zeros((B, feat.shape[1]), feat.dtype)])
        pool = np.arange(num_nodes, num_nodes + B)
        ei0 = np.repeat(pool, bnn)
        ei1 = np.arange(num_nodes)
        S.append(pool); D.append(pool); T.append(np.zeros(B, np.int64))
        if graph_pooling == "learninv":
            S.append(ei0); D.append(ei1); T.append(np.full(num_nodes, 3))
        S.append(ei1); D.append(ei0); T.append(np.full(num_nodes, 4))
        node_offset += B
        out["poolnodes"] = pool
    if conflicts == "node":
        M = max_n_machines
        machineid = feat[:num_nodes, 6].astype(np.int64)
        batchid = np.repeat(np.arange(B), bnn)
        feat = np.concatenate([feat, np.zeros((B * M, feat.shape[1]), feat.dtype)])
        mnodes = np.arange(node_offset, node_offset + M * B)
        idx = np.nonzero(machineid != -1)[0]
        mid = machineid[idx]
        target = batchid[idx] * M + mid + node_offset
        S.append(mnodes); D.append(mnodes); T.append(np.zeros(M * B, np.int64))
        S.append(idx); D.append(target); T.append(mid + 5)
        S.append(target); D.append(idx); T.append(mid + 5 + M)
        if graph_pooling == "learn":
            S.append(mnodes); D.append(np.repeat(np.arange(num_nodes, num_nodes + B), M))
            T.append(np.full(M * B, 5 + 2 * M))
        out["machinenodes"] = mnodes
    out.update(src=np.concatenate(S).astype(np.int64), dst=np.concatenate(D).astype(np.int64),
               type=np.concatenate(T).astype(np.int64), feat=feat, num_task_nodes=num_nodes)
    return out
