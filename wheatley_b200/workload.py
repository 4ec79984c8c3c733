"""Synthetic schedule graphs with the topology Wheatley's JSSP path produces.

Used by ``bench.py``, ``__graft_entry__.smoke()`` and the parity tests to make inputs of the
sizes BASELINE.json names without the reference tree (which does not exist on the GPU box).

Topology sources (reference): job chains ``jssp/env/state.py:220-227``; bidirectional
precedence edges typed 1/2 ``jssp/models/agent_observation.py:57-71``; machine cliques typed 5
``jssp/utils/utils.py:460-478`` + ``agent_observation.py:81-88``; self loops typed 0 ``:211``;
``dgl.batch`` ``:223``; pool nodes/edges ``jssp/models/gnn_dgl.py:263-286``.
Instances: machine order per job is a seeded random permutation and durations are U[1,99]
like ``instances/generate_taillard.py:19-36``."""
import numpy as np
import torch


def taillard_instance(n_j: int, n_m: int, seed: int):
    """(durations [n_j, n_m] int64 in [1, 99], machines [n_j, n_m] int64 permutation per job)."""
    rng = np.random.default_rng(seed)
    durations = rng.integers(1, 100, size=(n_j, n_m))
    machines = np.stack([rng.permutation(n_m) for _ in range(n_j)])
    return durations.astype(np.int64), machines.astype(np.int64)


def jssp_observation(n_j: int, n_m: int, seed: int, n_features: int = None, progress: float = 0.0):
    """One env observation in the layout of ``jssp/env/observation.py:83-128`` (unpadded):
    ``features [N, F]`` (col 6 = machine id, ``agent_observation.py:162``), ``edge_index [2, n_e]``
    = job-chain precedence edges, plus machine-chain edges between consecutively dispatched
    operations of a machine for a seeded random dispatch of ``progress`` of the operations
    (``jssp/env/state.py:655-674``)."""
    durations, machines = taillard_instance(n_j, n_m, seed)
    N = n_j * n_m
    F = n_features if n_features is not None else 6 + n_m + 4
    rng = np.random.default_rng(seed + 7919)
    feat = rng.random((N, F), dtype=np.float32)
    feat[:, 6] = machines.reshape(-1).astype(np.float32)
    feat[:, 4] = durations.reshape(-1).astype(np.float32) / 99.0
    ids = np.arange(N).reshape(n_j, n_m)
    src = [ids[:, :-1].reshape(-1)]
    dst = [ids[:, 1:].reshape(-1)]
    n_sched = int(round(progress * N))
    if n_sched > 0:
        nxt = np.zeros(n_j, dtype=np.int64)          # next operation of each job
        last = -np.ones(n_m, dtype=np.int64)         # last dispatched node per machine
        cs, cd = [], []
        for _ in range(n_sched):
            cand = np.nonzero(nxt < n_m)[0]
            j = int(cand[rng.integers(0, len(cand))])
            node = int(ids[j, nxt[j]])
            m = int(machines[j, nxt[j]])
            if last[m] >= 0 and ids.reshape(-1)[last[m]] // n_m != j:
                cs.append(last[m]); cd.append(node)
            last[m] = node
            nxt[j] += 1
        src.append(np.array(cs, dtype=np.int64)); dst.append(np.array(cd, dtype=np.int64))
    ei = np.stack([np.concatenate(src), np.concatenate(dst)]).astype(np.int64)
    return dict(features=feat, edge_index=ei, n_nodes=N, n_edges=ei.shape[1], n_jobs=n_j, n_machines=n_m)


def clique_edges(machineid: np.ndarray):
    """All ordered pairs (r, c), r != c, sharing a machine id (id -1 conflicts with nobody), in
    row-major order of r then c - the order ``jssp/utils/utils.py:460-478`` yields.  O(E) memory."""
    m = np.asarray(machineid).astype(np.int64)
    n = m.shape[0]
    valid = m >= 0
    nmach = int(m.max(initial=-1)) + 1
    if nmach <= 0:
        z = np.zeros(0, dtype=np.int64)
        return z, z, z
    order = np.argsort(np.where(valid, m, nmach), kind="stable")   # members per machine, ascending node id
    cnt = np.bincount(m[valid], minlength=nmach)
    start = np.concatenate([[0], np.cumsum(cnt)])
    nvalid = int(cnt.sum())
    rank = np.zeros(n, dtype=np.int64)                             # position of a node in its machine's list
    rank[order[:nvalid]] = np.arange(nvalid) - np.repeat(start[:-1], cnt)
    deg = np.where(valid, cnt[np.where(valid, m, 0)] - 1, 0)
    off = np.concatenate([[0], np.cumsum(deg)])
    r = np.repeat(np.arange(n), deg)
    k = np.arange(int(off[-1])) - off[r]                           # rank among r's neighbours
    kk = k + (k >= rank[r])                                        # skip r itself
    c = order[start[m[r]] + kk]
    return r.astype(np.int64), c.astype(np.int64), m[r]


def jssp_graph_edges(obs, conflicts="clique", mid_in_edges=False, bidir=True, add_self_loops=True):
    """Per-graph edge list in the reference order: [prec fwd (1) | prec rev (2)] | clique (5 or
    5+mid) | self loops (0).  ``agent_observation.py:53-88,186-211``."""
    e0, e1 = obs["edge_index"][0], obs["edge_index"][1]
    ne, nn = e0.shape[0], obs["n_nodes"]
    if bidir:
        s, d = np.concatenate([e0, e1]), np.concatenate([e1, e0])
        t = np.concatenate([np.full(ne, 1), np.full(ne, 2)])
    else:
        s, d, t = e0, e1, np.full(ne, 1)
    if conflicts == "clique":
        r, c, mid = clique_edges(obs["features"][:nn, 6].astype(np.int64))
        s, d = np.concatenate([s, r]), np.concatenate([d, c])
        t = np.concatenate([t, mid + 5 if mid_in_edges else np.full(r.shape[0], 5)])
    if add_self_loops:
        loop = np.arange(nn)
        s, d, t = np.concatenate([s, loop]), np.concatenate([d, loop]), np.concatenate([t, np.zeros(nn, np.int64)])
    return s.astype(np.int64), d.astype(np.int64), t.astype(np.int64)


def jssp_batch(n_j, n_m, batch, seed0=0, progress=0.0, conflicts="clique", pool=True, n_features=None):
    """A ``dgl.batch`` of ``batch`` schedule graphs + one learned-pool node per graph.
    Returns dict(src, dst, etype [E] int64 torch, feat [N, F] fp32, n_nodes, n_task_nodes,
    batch_num_nodes, edges_per_graph)."""
    S, D, T, feats, bnn = [], [], [], [], []
    off = 0
    for b in range(batch):
        obs = jssp_observation(n_j, n_m, seed0 + b, n_features, progress)
        s, d, t = jssp_graph_edges(obs, conflicts)
        S.append(s + off); D.append(d + off); T.append(t)
        feats.append(obs["features"]); bnn.append(obs["n_nodes"])
        off += obs["n_nodes"]
    n_task = off
    feat = np.concatenate(feats)
    if pool:                                               # jssp/models/gnn_dgl.py:263-286
        bnn_a = np.array(bnn)
        poolnodes = np.arange(n_task, n_task + batch)
        S.append(poolnodes); D.append(poolnodes); T.append(np.zeros(batch, np.int64))
        S.append(np.arange(n_task)); D.append(np.repeat(poolnodes, bnn_a)); T.append(np.full(n_task, 4))
        feat = np.concatenate([feat, np.zeros((batch, feat.shape[1]), np.float32)])
        off += batch
    src, dst, et = np.concatenate(S), np.concatenate(D), np.concatenate(T)
    return dict(src=torch.from_numpy(src), dst=torch.from_numpy(dst), etype=torch.from_numpy(et),
                feat=torch.from_numpy(feat), n_nodes=off, n_task_nodes=n_task,
                batch_num_nodes=bnn, n_edges=int(src.shape[0]))


def jssp_clique_graph(n_j, n_m, seed=0):
    """One full-size graph (tasks + pool node): ``(src, dst, etype, n_nodes)``."""
    b = jssp_batch(n_j, n_m, 1, seed)
    return b["src"], b["dst"], b["etype"], b["n_nodes"]


def jssp_observation_batch(n_j, n_m, batch, seed0=0, progress=0.0, n_features=None):
    """The padded gym observation of ``batch`` graphs, as ``AgentObservation.from_gym_observation`` receives it
    (jssp/models/agent_observation.py:127-131): ``features [B, max_nodes, F]``, ``edge_index [B, 2, max_edges]``
    (zero-padded), ``n_nodes [B]``, ``n_edges [B]`` - CPU torch tensors."""
    obs = [jssp_observation(n_j, n_m, seed0 + b, n_features, progress) for b in range(batch)]
    max_e = max(o["n_edges"] for o in obs)
    ei = np.zeros((batch, 2, max_e), np.int64)
    for i, o in enumerate(obs):
        ei[i, :, : o["n_edges"]] = o["edge_index"]
    return dict(features=torch.from_numpy(np.stack([o["features"] for o in obs])),
                edge_index=torch.from_numpy(ei),
                n_nodes=torch.tensor([o["n_nodes"] for o in obs], dtype=torch.int64),
                n_edges=torch.tensor([o["n_edges"] for o in obs], dtype=torch.int64))


# ------------------------------------------------------------------------- PSP / PyG flavour
def psp_taillard_graph(n_j: int, n_m: int, seed: int, n_rp: int = 0, progress: float = 0.0):
    """One env-side observation of a Taillard JSSP modelled as an RCPSP the way ``psp/utils/taillard_rcpsp.py:46-160``
    does (how every JSSP config runs today): node 0 = source, node 1 + j*n_m + i = operation i of job j,
    last node = sink; ``prec`` edges in ``problem_edges`` order (psp/env/gstate.py:98-103: by source node, the
    source's n_j successors first); one renewable resource per machine with capacity 1, so
    ``resources[node, machine] = 1`` (gstate.py:260-277).  ``feat [N, 16]``: the columns of
    agent_graph_observation.py:66-86 with seeded values (only the numerics depend on them); ``n_rp``
    resource-priority edges between operations of one machine with 4 attributes (gstate.py:891-934).
    Returns the dict ``wheatley_b200.wire.PspObsLayout.encode`` takes.  ``psp_taillard_graph(4, 4, s)`` has the
    shape of ``instances/psp/272/272.sm`` (18 jobs incl. source / sink, 4 resources)."""
    _, machines = taillard_instance(n_j, n_m, seed)
    N = n_j * n_m + 2
    ids = 1 + np.arange(n_j * n_m).reshape(n_j, n_m)
    nxt = np.concatenate([ids[:, 1:], np.full((n_j, 1), N - 1)], axis=1)
    prec = np.concatenate([np.stack([np.zeros(n_j, np.int64), ids[:, 0]]),
                           np.stack([ids.reshape(-1), nxt.reshape(-1)])], axis=1).astype(np.int64)
    resources = np.zeros((N, n_m), np.float32)
    resources[ids.reshape(-1), machines.reshape(-1)] = 1.0
    rng = np.random.default_rng(seed + 104729)
    feat = rng.random((N, 16), dtype=np.float32)
    n_aff = int(round(progress * N))
    feat[:, 0] = 0.0
    feat[:n_aff, 0] = 1.0                                  # affected
    feat[:, 1] = (rng.random(N) < 0.3).astype(np.float32)  # selectable
    feat[:, 2] = np.concatenate([[0], 1 + np.arange(n_j * n_m), [N - 1]]).astype(np.float32)   # job id
    feat[:, 3] = 0.0
    feat[0, 3], feat[-1, 3] = -1.0, 1.0                    # type: source / sink
    feat[:, 13] = 0.0                                      # has_due_date
    feat[:, 15] = (np.arange(N) < n_aff).astype(np.float32)   # past
    rp = np.zeros((2, 0), np.int64)
    rp_r = np.zeros((0, 4), np.float32)
    if n_rp:
        m = rng.integers(0, n_m, n_rp)
        flat_m = machines.reshape(-1)
        s, d = [], []
        for mm in m:
            users = ids.reshape(-1)[flat_m == mm]
            a, b = rng.choice(users, 2, replace=False)
            s.append(a); d.append(b)
        rp = np.array([s, d], dtype=np.int64)
        rp_r = np.concatenate([m[:, None].astype(np.float32), rng.random((n_rp, 3), dtype=np.float32)], axis=1)
    return dict(feat=feat, resources=resources, res_cal=np.zeros((n_m, 3), np.float32), prec=prec, rp=rp, rp_r=rp_r)


def psp_sizes_expected(n_j, n_m, conflicts="node", pooling=True):
    """(nodes, edges) per graph at t = 0, SURVEY 8d table (PyG flavour)."""
    N = n_j * n_m + 2
    prec = n_j * (n_m + 1)
    e = 2 * prec + N
    n = N
    if conflicts == "node":
        n += n_m
        e += 2 * n_j * n_m
    elif conflicts == "clique":
        e += n_m * n_j * (n_j - 1)
    if pooling:
        n += 1
        e += N + (n_m if conflicts == "node" else 0) + 1
    return n, e
