"""numpy restatement of the PSP / PyG-flavour batch construction:
``Agent._obs_as_tensor_graph`` -> ``AgentGraphObservation.rewire_internal`` (psp/models/agent.py:257-270,
psp/models/agent_graph_observation.py:55-118) followed by ``Agent.preprocess`` (psp/models/agent.py:298-406):
``Batch.from_data_list`` per edge type, self loops, ``rewire`` (resource nodes, pool nodes and their edges,
psp/models/rewiring.py:27-363), ``homogeneous_edges`` (:366-440) and ``to_homogeneous``.
TEST INFRASTRUCTURE - see oracle/__init__.py.  Pinned by tests/golden/psp_batch.npz: the reference functions
above executed verbatim on oracle/shims/psp_graph.py (tests/golden/make_golden.py::golden_psp_batch).

Integer work and copied fp32 attributes: bit-exact.

Per-graph input (what an env worker's graph holds, psp/env/gstate.py): ``feat [n, 16]`` (the 16 columns
``rewire_internal`` concatenates, agent_graph_observation.py:66-86), ``resources [n, R]``, ``res_cal [R, 3]``,
``prec [2, P]``, ``rp [2, Q]``, ``rp_r [Q, 4]``."""
import numpy as np

EDGE_TYPES = {"self": 0, "prec": 1, "rprec": 2, "rc": 3, "rp": 4, "rrp": 5, "pool": 6, "rpool": 7, "vnode": 8,
              "rvnode": 9, "nodeconf": 10, "rnodeconf": 11, "poolres": 12, "rpoolres": 13, "selfpool": 14,
              "selfres": 15}                                   # agent_graph_observation.py:36-53


def base_store_order(conflicts, first_graph_env_stores=("prec",)):
    """Order in which the FIRST graph of the batch got its edge stores (the batch inherits it,
    ``Batch.from_data_list``): the env's own stores (``prec``, then ``rc`` when the env observes conflicts as
    cliques, ``rp`` once it has added a resource-priority edge), then ``rewire_internal``
    (agent_graph_observation.py:88-118): ``rc`` (clique mode, on the ``num_edges`` probe :89), ``rprec`` (:108),
    ``rp`` (probe :110), ``rrp`` (:112 or fill_void)."""
    order = list(first_graph_env_stores)
    for t in (["rc"] if conflicts == "clique" else []) + ["rprec", "rp", "rrp"]:
        if t not in order:
            order.append(t)
    return order


def resource_conflicts(r_info):
    """psp/utils/utils.py:73-108 ``compute_resources_graph_torch``: every ordered pair (i, j), i != j, and
    resource r that both use, in ``torch.where`` order of the [i, j, r] mask; returns
    ``(src=i, dst=j, rid=r, val=r_info[i, r], valr=r_info[j, r])``."""
    r_info = np.asarray(r_info)
    used = r_info != 0
    c = used[:, None, :] & used[None, :, :] & ~np.eye(r_info.shape[0], dtype=bool)[:, :, None]
    i, j, r = np.nonzero(c)
    return i.astype(np.int64), j.astype(np.int64), r.astype(np.int64), r_info[i, r], r_info[j, r]


def psp_batch(graphs, conflicts="clique", graph_pooling="learn", self_loops=True, store_order=None):
    """Returns dict(edge_index [2, E] int64, type int64 [E], rid int32 [E], att_rc fp32 [E, 2],
    att_rp fp32 [E, 3], feat fp32 [N_total, 16] (resource-node rows: zeros here, uninitialised in the reference,
    rewiring.py:132-141), batch_size, num_nodes, n_nodes, poolnodes, resource_nodes, res_cal_id,
    segments = [(edge type name, first edge, count)])."""
    B = len(graphs)
    if store_order is None:
        store_order = base_store_order(conflicts)
    bnn = np.array([g["feat"].shape[0] for g in graphs], dtype=np.int64)
    off = np.concatenate([[0], np.cumsum(bnn)])
    NT = int(off[-1])
    Rmax = max(g["resources"].shape[1] for g in graphs)               # pad_resources, pyg_graph.py:203-215
    res = np.concatenate([np.pad(g["resources"], ((0, 0), (0, Rmax - g["resources"].shape[1]))) for g in graphs])
    res_cal_id = np.concatenate([np.pad(g["res_cal"], ((0, Rmax - g["res_cal"].shape[0]), (0, 0)))
                                 for g in graphs]).astype(np.float32)  # agent_graph_observation.py:264-269
    S, D, T, RID, ARC, ARP, seg = [], [], [], [], [], [], []

    def emit(name, s, d, rid=None, att_rc=None, att_rp=None):
        n = len(s)
        seg.append((name, sum(len(x) for x in S), n))
        S.append(np.asarray(s, np.int64)); D.append(np.asarray(d, np.int64))
        T.append(np.full(n, EDGE_TYPES[name], np.int64))
        RID.append(np.zeros(n, np.int32) if rid is None else np.asarray(rid).astype(np.int32))
        ARC.append(np.zeros((n, 2), np.float32) if att_rc is None else np.asarray(att_rc, np.float32))
        ARP.append(np.zeros((n, 3), np.float32) if att_rp is None else np.asarray(att_rp, np.float32))

    def rp_attrs(r):                                                    # rewiring.py:411-429
        a = np.array(r[:, 1:], dtype=np.float32, copy=True)
        a[:, -1] += 1
        return (r[:, 0] + 1).astype(np.int32), a

    for t in store_order:                                               # Batch.from_data_list, per store
        if t in ("prec", "rprec"):
            e = [g["prec"] + o for g, o in zip(graphs, off)]
            e = np.concatenate(e, axis=1) if e else np.zeros((2, 0), np.int64)
            emit(t, e[0], e[1]) if t == "prec" else emit(t, e[1], e[0])           # :107-109
        elif t in ("rp", "rrp"):
            e = np.concatenate([g["rp"] + o for g, o in zip(graphs, off)], axis=1)
            r = np.concatenate([g["rp_r"].reshape(-1, 4) for g in graphs])
            rid, a = rp_attrs(r) if r.shape[0] else (None, None)
            emit(t, e[0], e[1], rid, None, a) if t == "rp" else emit(t, e[1], e[0], rid, None, a)   # :111-118
        elif t == "rc" and conflicts == "clique":                        # :88-105, rewiring.py:431-438
            parts = [resource_conflicts(g["resources"]) for g in graphs]
            s = np.concatenate([p[0] + o for p, o in zip(parts, off)])
            d = np.concatenate([p[1] + o for p, o in zip(parts, off)])
            rid = np.concatenate([p[2] for p in parts]) + 1
            arc = np.stack([np.concatenate([p[3] for p in parts]), np.concatenate([p[4] for p in parts])], axis=1)
            emit("rc", s, d, rid, arc)
    if self_loops:                                                      # agent.py:339-344, pyg_graph.py:331-335
        emit("self", np.arange(NT), np.arange(NT))
    node_offset = NT
    n_extra = 0
    resource_nodes = np.zeros(0, np.int64)
    if conflicts == "node":                                             # rewiring.py:104-253
        resource_nodes = np.arange(node_offset, node_offset + Rmax * B)
        batch_id = np.repeat(np.arange(B), bnn)
        consumers, ridx = np.nonzero(res != 0)
        resource_index = ridx + (node_offset + Rmax * batch_id[consumers])
        rc = np.repeat(res[consumers, ridx][:, None], 2, axis=1)
        emit("nodeconf", consumers, resource_index, ridx, rc)
        emit("rnodeconf", resource_index, consumers, ridx, rc)
        # :218-228 - indexes the [B * Rmax] node list with COLUMN ids of resources unused by the whole batch
        unused = resource_nodes[np.nonzero(np.all(res == 0, axis=0))[0]]
        emit("selfres", unused, unused)
        node_offset += Rmax * B
        n_extra += Rmax * B
    poolnodes = np.zeros(0, np.int64)
    if graph_pooling in ("learn", "learninv"):                          # rewiring.py:27-101
        poolnodes = np.arange(node_offset, node_offset + B)
        ei0 = np.repeat(poolnodes, bnn)
        ei1 = np.arange(NT)
        emit("pool", ei1, ei0)
        if graph_pooling == "learninv":
            emit("rpool", ei0, ei1)
        if conflicts == "node":
            emit("poolres", resource_nodes, np.repeat(poolnodes, Rmax))
            if graph_pooling == "learninv":
                emit("rpoolres", np.repeat(poolnodes, Rmax), resource_nodes)
        emit("selfpool", poolnodes, poolnodes)
        node_offset += B
        n_extra += B
    feat = np.concatenate([np.concatenate([g["feat"] for g in graphs]).astype(np.float32),
                           np.zeros((n_extra, 16), np.float32)])
    return dict(edge_index=np.stack([np.concatenate(S), np.concatenate(D)]), type=np.concatenate(T),
                rid=np.concatenate(RID), att_rc=np.concatenate(ARC), att_rp=np.concatenate(ARP), feat=feat,
                batch_size=B, num_nodes=NT, n_nodes=bnn, poolnodes=poolnodes, resource_nodes=resource_nodes,
                res_cal_id=res_cal_id, segments=seg, n_total=node_offset, max_n_resources=Rmax)
