"""Device-side JSSP batch builder - integer / byte work, bit-exact:
against the golden vectors produced by the reference's own ``AgentObservation.from_gym_observation``
(tests/golden/jssp_batch.npz), against the numpy oracle for the option space and the pool wiring, and at the
bench size against the host builder of the workload generator."""
import os
import zlib

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def wb():
    import wheatley_b200
    return wheatley_b200


def _obs_t(features, edge_index, n_nodes, n_edges):
    return dict(features=torch.from_numpy(np.ascontiguousarray(features)),
                edge_index=torch.from_numpy(np.ascontiguousarray(edge_index)),
                n_nodes=torch.as_tensor(np.asarray(n_nodes)), n_edges=torch.as_tensor(np.asarray(n_edges)))


def _same(dev, ref_src, ref_dst, ref_type, ref_feat):
    assert np.array_equal(dev["src"].cpu().numpy(), ref_src)
    assert np.array_equal(dev["dst"].cpu().numpy(), ref_dst)
    assert np.array_equal(dev["etype"].cpu().numpy(), ref_type)
    assert np.array_equal(dev["feat"].cpu().numpy(), ref_feat)          # copied / selected floats: exact


@pytest.mark.parametrize("tag,conflicts,mie", [("att", "att", False), ("clique", "clique", False),
                                               ("cliquemid", "clique", True)])
def test_batch_matches_reference_golden(wb, golden_dir, tag, conflicts, mie):
    from wheatley_b200.batch import jssp_batch_on_device
    g = np.load(os.path.join(str(golden_dir), "jssp_batch.npz"))
    obs = _obs_t(g[f"{tag}.obs.features"], g[f"{tag}.obs.edge_index"], g[f"{tag}.obs.n_nodes"],
                 g[f"{tag}.obs.n_edges"])
    dev = jssp_batch_on_device(obs, "cuda:0", conflicts=conflicts, graph_pooling=None, mid_in_edges=mie)
    _same(dev, g[f"{tag}.src"], g[f"{tag}.dst"], g[f"{tag}.type"], g[f"{tag}.feat"])
    assert dev["batch_num_nodes"] == g[f"{tag}.bnn"].tolist() and dev["poolnodes"] is None


def _random_obs(rng, B, max_nodes, F, n_mach, p_unaffected=0.2):
    nn = rng.integers(1, max_nodes + 1, B)
    nn[rng.integers(0, B)] = max_nodes
    ne = np.array([rng.integers(0, 2 * n + 1) for n in nn])
    feat = rng.standard_normal((B, max_nodes, F)).astype(np.float32)
    mid = rng.integers(0, n_mach, (B, max_nodes)).astype(np.float32)
    mid[rng.random((B, max_nodes)) < p_unaffected] = -1
    feat[:, :, 6] = mid
    feat[:, :, 1] = np.where(rng.random((B, max_nodes)) < 0.5, -1, feat[:, :, 1])
    ei = np.zeros((B, 2, max(int(ne.max()), 1)), np.int64)
    for i in range(B):
        ei[i, :, : ne[i]] = rng.integers(0, nn[i], (2, ne[i]))
    return feat, ei, nn.astype(np.int64), ne.astype(np.int64)


@pytest.mark.parametrize("conflicts", ["att", "clique"])
@pytest.mark.parametrize("pooling", [None, "learn", "learninv"])
@pytest.mark.parametrize("opts", [dict(), dict(bidir=False), dict(add_self_loops=False),
                                  dict(no_tct=True, mid_in_edges=True)])
def test_batch_matches_oracle_option_space(wb, conflicts, pooling, opts):
    from oracle import batch_oracle
    from wheatley_b200.batch import jssp_batch_on_device
    rng = np.random.default_rng(zlib.crc32(repr((conflicts, pooling, sorted(opts))).encode()))
    feat, ei, nn, ne = _random_obs(rng, B=5, max_nodes=77, F=11, n_mach=6)
    ref = batch_oracle.jssp_batch(feat, ei, nn, ne, conflicts=conflicts, **opts)
    if pooling:
        ref = batch_oracle.jssp_add_pool_and_machine_nodes(ref, pooling, conflicts)
    dev = jssp_batch_on_device(_obs_t(feat, ei, nn, ne), "cuda:0", conflicts=conflicts, graph_pooling=pooling,
                               **opts)
    _same(dev, ref["src"], ref["dst"], ref["type"], ref["feat"])
    if pooling:
        assert np.array_equal(dev["poolnodes"].cpu().numpy(), ref["poolnodes"])
        assert dev["n_task_nodes"] == ref["num_task_nodes"]


@pytest.mark.parametrize("pooling", [None, "learn", "learninv"])
@pytest.mark.parametrize("M", [6, 9])
def test_batch_node_conflicts_match_oracle(wb, pooling, M):
    """conflicts="node" (jssp/models/gnn_dgl.py:288-335): machine nodes after the pool nodes, self loops, task ->
    machine, machine -> task (ascending affected task order), machine -> pool for "learn"."""
    from oracle import batch_oracle
    from wheatley_b200.batch import jssp_batch_on_device
    rng = np.random.default_rng(100 + M + len(str(pooling)))
    feat, ei, nn, ne = _random_obs(rng, B=4, max_nodes=70, F=9, n_mach=6, p_unaffected=0.3)
    ref = batch_oracle.jssp_batch(feat, ei, nn, ne, conflicts="node")
    ref = batch_oracle.jssp_add_pool_and_machine_nodes(ref, pooling, "node", max_n_machines=M)
    dev = jssp_batch_on_device(_obs_t(feat, ei, nn, ne), "cuda:0", conflicts="node", graph_pooling=pooling,
                               max_n_machines=M)
    _same(dev, ref["src"], ref["dst"], ref["type"], ref["feat"])
    assert np.array_equal(dev["machinenodes"].cpu().numpy(), ref["machinenodes"])
    assert dev["n_nodes"] == ref["feat"].shape[0]
    with pytest.raises(wb.B200GatError):
        jssp_batch_on_device(_obs_t(feat, ei, nn, ne), "cuda:0", conflicts="node", graph_pooling=pooling,
                             max_n_machines=3)


def test_batch_edge_cases(wb):
    from oracle import batch_oracle
    from wheatley_b200.batch import jssp_batch_on_device
    # one graph, one node, no precedence edge, unaffected machine
    feat = np.zeros((1, 1, 8), np.float32)
    feat[0, 0, 6] = -1
    ei = np.zeros((1, 2, 0), np.int64)
    ref = batch_oracle.jssp_add_pool_and_machine_nodes(
        batch_oracle.jssp_batch(feat, ei, [1], [0], conflicts="clique"), "learn", "clique")
    dev = jssp_batch_on_device(_obs_t(feat, ei, [1], [0]), "cuda:0", conflicts="clique", graph_pooling="learn")
    _same(dev, ref["src"], ref["dst"], ref["type"], ref["feat"])
    # every node on one machine: a single full clique of 300 nodes (more rows than one CTA handles)
    rng = np.random.default_rng(3)
    feat = rng.standard_normal((2, 300, 9)).astype(np.float32)
    feat[:, :, 6] = 0
    ei = rng.integers(0, 300, (2, 2, 40))
    ref = batch_oracle.jssp_batch(feat, ei, [300, 299], [40, 7], conflicts="clique")
    dev = jssp_batch_on_device(_obs_t(feat, ei, [300, 299], [40, 7]), "cuda:0", conflicts="clique",
                               graph_pooling=None)
    _same(dev, ref["src"], ref["dst"], ref["type"], ref["feat"])
    assert dev["n_edges"] == 2 * 47 + 300 * 299 + 299 * 298 + 599
    with pytest.raises(ValueError):
        jssp_batch_on_device(_obs_t(feat, ei, [300, 299], [40, 7]), "cuda:0", conflicts="node")     # no max_n_machines
    with pytest.raises(wb.B200GatError):
        jssp_batch_on_device(_obs_t(feat, ei, [301, 299], [40, 7]), "cuda:0", conflicts="clique")


def test_batch_bench_size_and_trunk(wb):
    """8 x Taillard 100x20 (1.65 M edges): the device-built batch equals the host-built one bit for bit, and the
    trunk gives the same output on it (same plan -> same kernels -> same bits)."""
    from wheatley_b200.agent import JsspAgent
    from wheatley_b200.batch import jssp_batch_from_observation
    from wheatley_b200.gnn_dgl import JsspBatch
    from wheatley_b200.workload import jssp_batch, jssp_observation_batch
    n_j, n_m, B = 100, 20, 8
    hb = jssp_batch(n_j, n_m, B, seed0=0, progress=0.3)
    obs = jssp_observation_batch(n_j, n_m, B, seed0=0, progress=0.3)
    jb, dev = jssp_batch_from_observation(obs, "cuda:0", n_j * n_m, conflicts="clique", graph_pooling="learn")
    feat_ref = hb["feat"].numpy().copy()            # the generator's raw features + the builder's fix-up (:133-135)
    feat_ref[:, 1] = np.where(feat_ref[:, 1] == -1, -1, feat_ref[:, 4])
    feat_ref[hb["n_task_nodes"]:] = 0
    _same(dev, hb["src"].numpy(), hb["dst"].numpy(), hb["etype"].numpy(), feat_ref)
    assert dev["n_edges"] == hb["n_edges"] and dev["n_nodes"] == hb["n_nodes"]
    torch.manual_seed(0)
    agent = JsspAgent(hb["feat"].shape[1], n_j * n_m, n_m, 64, 2, 4).cuda()
    NT = hb["n_task_nodes"]
    plan = wb.GraphPlan(hb["src"].cuda(), hb["dst"].cuda(), hb["n_nodes"])
    jb_host = JsspBatch(plan, hb["etype"].cuda(), torch.from_numpy(feat_ref).cuda(), B, n_j * n_m, NT,
                        poolnodes=torch.arange(NT, NT + B).cuda())
    with torch.no_grad():
        a = agent.gnn.node_embeddings(jb_host)
        b = agent.gnn.node_embeddings(jb)
    assert torch.equal(a, b)


@pytest.mark.parametrize("side_stream", [False, True])
def test_prefetcher_observation_path(wb, side_stream):
    from wheatley_b200.batch import jssp_batch_from_observation
    from wheatley_b200.prefetch import BatchPrefetcher
    from wheatley_b200.workload import jssp_observation_batch
    n_j, n_m, B = 12, 5, 3
    obs = {k: v.pin_memory() for k, v in jssp_observation_batch(n_j, n_m, B, seed0=4, progress=0.4).items()}
    extras = dict(adv=torch.randn(B).pin_memory())
    pf = BatchPrefetcher("cuda:0", side_stream=side_stream)
    pf.submit_observation(obs, n_j * n_m, extras, conflicts="clique", graph_pooling="learn")
    d, plan = pf.get()
    jb, ref = jssp_batch_from_observation(obs, "cuda:0", n_j * n_m, conflicts="clique", graph_pooling="learn")
    torch.cuda.synchronize()
    assert d["batch"].plan is plan and torch.equal(d["adv"].cpu(), extras["adv"])
    for k in ("src", "dst", "etype", "feat"):
        assert torch.equal(d[k], ref[k]), k
    va, vb = next(iter(plan._clique_views.values())), next(iter(jb.plan._clique_views.values()))
    assert va._pending is None and va.n_blocks == vb.n_blocks > 0 and va.n_res_edges == vb.n_res_edges
