"""GPU parity of the PSP / PyG-flavour batch builder: ``b200gat_psp_batch_build`` over wire-format
observation slots against (i) the golden vectors produced by executing the reference's own ``Agent.preprocess``
(psp/models/agent.py:298-406 + rewiring.py + agent_graph_observation.py) and (ii) the numpy oracle at the
BASELINE.json sizes.  Integer work and copied attributes: bit-exact, same edge order."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

PSP_TAGS = ["node272", "clique272", "nodeinv", "ragged", "clique_noself", "single"]


@pytest.fixture(scope="module")
def wb():
    import wheatley_b200
    return wheatley_b200


def golden_case(g, tag):
    conflicts, pooling, _vnode, self_loops = (str(x) for x in g[f"{tag}.opts"])
    B = int(g[f"{tag}.n_graphs"])
    graphs = [dict(feat=g[f"{tag}.g{i}.feat"], resources=g[f"{tag}.g{i}.resources"],
                   res_cal=g[f"{tag}.g{i}.res_cal"], prec=g[f"{tag}.g{i}.prec"], rp=g[f"{tag}.g{i}.rp"],
                   rp_r=g[f"{tag}.g{i}.rp_r"]) for i in range(B)]
    return graphs, conflicts, pooling, self_loops == "1", [str(x) for x in g[f"{tag}.g0.stores_env"]]


def compare(d, ref, resource_nodes):
    """``d``: psp_batch_on_device output; ``ref``: dict of numpy arrays with the reference's names."""
    torch.cuda.synchronize()
    got = {"edge_index": d["edge_index"], "type": d["type"], "rid": d["rid"], "att_rc": d["attr"][:, :2],
           "att_rp": d["attr"][:, 2:], "res_cal_id": d["res_cal_id"]}
    for k, v in got.items():
        v = v.cpu().numpy()
        assert v.dtype == ref[k].dtype and v.shape == ref[k].shape, (k, v.dtype, v.shape, ref[k].dtype, ref[k].shape)
        assert np.array_equal(v, ref[k]), f"{k} differs from the reference"
    feat = d["feat"].cpu().numpy()
    assert feat.shape == ref["feat"].shape
    keep = np.ones(feat.shape[0], bool)
    keep[resource_nodes] = False                        # torch.empty rows in the reference (rewiring.py:132-141)
    assert np.array_equal(feat[keep], ref["feat"][keep])


@pytest.mark.parametrize("tag", PSP_TAGS)
def test_psp_batch_matches_reference_golden(wb, golden_dir, tag):
    from wheatley_b200.psp_batch import preprocess, psp_batch_on_device
    from wheatley_b200.wire import PspArena
    g = np.load(os.path.join(golden_dir, "psp_batch.npz"))
    graphs, conflicts, pooling, self_loops, env_stores = golden_case(g, tag)
    arena = PspArena.from_graphs(graphs)
    d = psp_batch_on_device(arena, "cuda", conflicts=conflicts, graph_pooling=pooling, self_loops=self_loops,
                            first_graph_env_stores=env_stores)
    ref = {k: g[f"{tag}.out.{k}"] for k in ("edge_index", "type", "rid", "att_rc", "att_rp", "res_cal_id", "feat")}
    rn = g[f"{tag}.out.resource_nodes"]
    compare(d, ref, rn)
    assert d["batch_size"] == int(g[f"{tag}.out.batch_size"]) and d["num_nodes"] == int(g[f"{tag}.out.num_nodes"])
    assert d["batch_num_nodes"] == g[f"{tag}.out.n_nodes"].tolist()
    assert (d["poolnodes"] or []) == g[f"{tag}.out.poolnodes"].tolist()
    assert ([] if d["resource_nodes"] is None else d["resource_nodes"].tolist()) == rn.tolist()
    # the 9-tuple of Agent.preprocess (agent.py:396-406)
    tup = preprocess(arena, "cuda", conflicts=conflicts, graph_pooling=pooling, self_loops=self_loops,
                     first_graph_env_stores=env_stores)
    assert len(tup) == 9 and tup[6] is None and tup[8] is None
    assert tup[1] == d["batch_size"] and tup[2] == d["num_nodes"]
    n_nodes = g[f"{tag}.out.n_nodes"].tolist()
    assert tup[3] == (n_nodes[0] if len(n_nodes) == 1 else n_nodes)      # an int for a batch of one (:315-319)
    assert torch.equal(tup[0].edata(None, "type"), d["type"]) and tup[0].plan.n_edges == d["n_edges"]


@pytest.mark.parametrize("n_j,n_m,B,conflicts,pooling", [(4, 4, 64, "node", "learn"), (10, 10, 5, "clique", "learn"),
                                                         (15, 15, 3, "node", "learninv"), (30, 20, 2, "clique", "max"),
                                                         (100, 20, 2, "node", "learn"), (100, 20, 1, "clique", "learn")])
def test_psp_batch_matches_oracle_at_baseline_sizes(wb, n_j, n_m, B, conflicts, pooling):
    from oracle import psp_batch_oracle as pbo
    from wheatley_b200.psp_batch import psp_batch_on_device
    from wheatley_b200.wire import PspArena
    from wheatley_b200.workload import psp_sizes_expected, psp_taillard_graph
    graphs = [psp_taillard_graph(n_j, n_m, 7 + b, n_rp=(3 * b) % 7, progress=0.1 * b) for b in range(B)]
    arena = PspArena.from_graphs(graphs)
    d = psp_batch_on_device(arena, "cuda", conflicts=conflicts, graph_pooling=pooling)
    ref = pbo.psp_batch(graphs, conflicts, pooling, True)
    compare(d, ref, ref["resource_nodes"])
    n_exp, e_exp = psp_sizes_expected(n_j, n_m, conflicts, pooling in ("learn", "learninv"))
    n_rp_total = sum(g["rp"].shape[1] for g in graphs)
    extra = (n_j * n_m + 2 + (n_m if conflicts == "node" else 0)) * B if pooling == "learninv" else 0
    assert d["n_total"] == n_exp * B and d["n_edges"] == e_exp * B + 2 * n_rp_total + extra


def test_psp_batch_store_orders_and_edge_cases(wb):
    from oracle import psp_batch_oracle as pbo
    from wheatley_b200.psp_batch import psp_batch_on_device
    from wheatley_b200.wire import PspArena, PspObsLayout
    from wheatley_b200.workload import psp_taillard_graph
    graphs = [psp_taillard_graph(5, 3, 1, n_rp=4), psp_taillard_graph(3, 3, 2), psp_taillard_graph(6, 2, 3, n_rp=1)]
    # a resource nobody uses (selfres edge), a layout with slack, every store order the env can produce
    for g_ in graphs:
        g_["resources"] = np.concatenate([g_["resources"], np.zeros((g_["resources"].shape[0], 2), np.float32)], 1)
        g_["res_cal"] = np.concatenate([g_["res_cal"], np.ones((2, 3), np.float32)])
    arena = PspArena.from_graphs(graphs, PspObsLayout(40, 7, 50, 9))
    for conflicts in ("node", "clique", "att"):
        for env_stores in (("prec",), ("prec", "rp"), ("prec", "rc", "rp"), ("prec", "rc")):
            order = pbo.base_store_order(conflicts, env_stores)
            d = psp_batch_on_device(arena, "cuda", conflicts=conflicts, graph_pooling="learn",
                                    first_graph_env_stores=env_stores)
            ref = pbo.psp_batch(graphs, conflicts, "learn", True, store_order=order)
            compare(d, ref, ref["resource_nodes"])
    with pytest.raises(TypeError):
        psp_batch_on_device(arena, "cuda", conflicts="node", vnode=True)
    with pytest.raises(wb.B200GatError):
        psp_batch_on_device(arena, "cpu")
