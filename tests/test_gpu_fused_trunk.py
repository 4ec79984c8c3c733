"""wheatley_b200.trunk (b200gat_trunk_fwd/bwd: all L layers behind one autograd node, one C-ABI call per pass)
against the layer-by-layer host mirror (``WHEATLEY_B200_TRUNK=layers``), which the other GPU tests pin to the
reference goldens / the fp64 oracle.  Both paths launch the same kernels on the same operands, so forward values must
be bit-identical; gradients differ only by where the residual / readout gradient is added (GEMM epilogue vs a
separate add): see ``grad_tol``.  Run in both arithmetic modes of the dense path (TF32 and 3xTF32)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

T = lambda a: torch.from_numpy(np.asarray(a))


@pytest.fixture(scope="module")
def wb():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import wheatley_b200
    return wheatley_b200


@pytest.fixture(params=[True, False], ids=["tf32", "fp32x3"])
def tf32(request):
    """Both arithmetic modes of the tcgen05 GEMM (csrc/gemm_tf32.cu): TF32 and 3xTF32 (fp32 accuracy)."""
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = request.param
    yield request.param
    torch.backends.cuda.matmul.allow_tf32 = prev
    os.environ.pop("WHEATLEY_B200_TRUNK", None)


def rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max()) / max(float(b.abs().max()), 1e-6)


ATOMIC = ("attn", ".att", "fc_fij", "lin_edge", "edge_embedder", "bias")


def grad_tol(name, tf32):
    """Both paths launch the same kernels; they add the residual / readout gradients in a different order, i.e. the
    operands of the following GEMMs differ in the last fp32 bit.  3xTF32 (fp32-accurate) GEMMs: 1e-5 (measured
    1.4e-6), 3e-4 for the gradients the attention backward accumulates with atomics (attention vector, edge-type
    table / affine map and what is upstream of them: not bit-reproducible run to run).  TF32 GEMMs keep 10 mantissa
    bits of each operand, so a last-bit difference can move an operand by 2^-11 = 4.9e-4 of its value: 1e-3
    (measured 2.3e-5 ... 3.5e-4 depending on which elements straddle a truncation boundary)."""
    if tf32:
        return 1e-3
    return 3e-4 if any(t in name for t in ATOMIC) else 1e-5


def _run_both(make_inputs, module, loss_of):
    outs = {}
    for mode in ("layers", "fused"):
        os.environ["WHEATLEY_B200_TRUNK"] = mode
        for p in module.parameters():
            p.grad = None
        from wheatley_b200 import _lib
        n0 = _lib.launch_count()
        res = module(make_inputs())
        loss_of(res).backward()
        torch.cuda.synchronize()
        outs[mode] = ([r.detach().clone() for r in (res if isinstance(res, (tuple, list)) else [res])],
                      {k: p.grad.detach().clone() for k, p in module.named_parameters() if p.grad is not None},
                      _lib.launch_count() - n0)
    return outs


@pytest.mark.parametrize("conflicts,pooling", [("clique", "learn"), ("node", "learn"), ("clique", "max")])
def test_gnn_dgl_fused_trunk_equals_layer_path(wb, tf32, conflicts, pooling):
    from wheatley_b200.batch import jssp_batch_from_observation
    from wheatley_b200.gnn_dgl import GnnDGL
    from wheatley_b200.workload import jssp_observation_batch
    n_j, n_m, B, L, Hd, H = 12, 5, 3, 3, 64, 4
    obs = jssp_observation_batch(n_j, n_m, B, seed0=4, progress=0.4)
    Fdim = obs["features"].shape[2]
    torch.manual_seed(2)
    mie = conflicts == "node"                # machine-node edge types are mid + 5 (+ M): the table needs them
    gnn = GnnDGL(Fdim, "gatv2", pooling, False, n_j * n_m, n_m, 3, L, Hd, "gelu", H, residual=True, normalize=False,
                 conflicts=conflicts, mid_in_edges=mie, split_output=True).cuda()

    def inputs():
        jb, _ = jssp_batch_from_observation(obs, "cuda", n_j * n_m, conflicts=conflicts, graph_pooling=pooling,
                                            mid_in_edges=mie, max_n_machines=n_m if conflicts == "node" else 0)
        return jb

    gen = torch.Generator().manual_seed(0)
    R = None

    def loss_of(res):
        nonlocal R
        nf, ge = res
        if R is None:
            R = (torch.randn(nf.shape, generator=gen).cuda(), torch.randn(ge.shape, generator=gen).cuda())
        return (nf * R[0]).sum() + (ge * R[1]).sum()

    o = _run_both(inputs, gnn, loss_of)
    for a, b in zip(o["fused"][0], o["layers"][0]):
        assert torch.equal(a, b)
    assert set(o["fused"][1]) == set(o["layers"][1])
    for k, g in o["fused"][1].items():
        assert rel(g, o["layers"][1][k]) < grad_tol(k, tf32), k
    assert o["fused"][2] >= o["layers"][2] - 2 * L          # same kernels (only the gradient adds moved)


@pytest.mark.parametrize("tag", ["node272", "clique272", "ragged_max"])
def test_gnn_mp_fused_trunk_equals_layer_path(wb, golden_dir, tf32, tag):
    """PSP flavour (GATv2 layers with lin_l / lin_r biases, output bias, EDGE_AFFINE edge features) on the inputs of
    the reference-generated trunk golden; also against that golden at the TF32 tolerance."""
    from wheatley_b200.gnn_mp import GnnMP
    from wheatley_b200.psp_batch import preprocess
    from wheatley_b200.wire import PspArena
    g = np.load(os.path.join(golden_dir, "gnn_mp.npz"))
    conflicts, pooling, _v, _s = (str(x) for x in g[f"{tag}.opts"])
    B = int(g[f"{tag}.n_graphs"])
    graphs = [dict(feat=g[f"{tag}.g{i}.feat"], resources=g[f"{tag}.g{i}.resources"],
                   res_cal=g[f"{tag}.g{i}.res_cal"], prec=g[f"{tag}.g{i}.prec"], rp=g[f"{tag}.g{i}.rp"],
                   rp_r=g[f"{tag}.g{i}.rp_r"]) for i in range(B)]
    env_stores = [str(x) for x in g[f"{tag}.g0.stores_env"]]
    Hd, H, L, max_n_nodes, max_res = (int(v) for v in g[f"{tag}.cfg"])
    gnn = GnnMP(16, pooling, max_n_nodes, max_res, 3, L, Hd, "gelu", H, residual=True, normalize=False,
                conflicts=conflicts, layer_pooling="all", add_rp_edges="frontier_strict", rwpe_k=0, rwpe_h=0,
                graphobs=True, shared_conv=False, pyg=True).cuda()
    gnn.load_state_dict({k[len(tag) + 3:]: T(g[k]) for k in g.files if k.startswith(tag + ".p.")})
    R1, R2 = T(g[f"{tag}.R1"]).cuda(), T(g[f"{tag}.R2"]).cuda()

    def inputs():
        return preprocess(PspArena.from_graphs(graphs), "cuda", conflicts=conflicts, graph_pooling=pooling,
                          first_graph_env_stores=env_stores)

    o = _run_both(inputs, gnn, lambda res: (res[0] * R1).sum() + (res[1] * R2).sum())
    for a, b in zip(o["fused"][0], o["layers"][0]):
        assert torch.equal(a, b)
    for k, gr in o["fused"][1].items():
        assert rel(gr, o["layers"][1][k]) < grad_tol(k, tf32), k
    # and against the reference's own GnnMP.forward (fp32) at the TF32 bound of DESIGN.md section 2
    assert rel(o["fused"][0][0], T(g[f"{tag}.node_features"])) < 2e-3
    assert rel(o["fused"][0][1], T(g[f"{tag}.graph_embedding"])) < 2e-3
    for k, gr in o["fused"][1].items():
        assert rel(gr, T(g[f"{tag}.g.{k}"])) < 5e-2, k          # measured worst: 2.7e-2 (rc_att_embedder, clique272)


def test_fused_trunk_inference_and_launch_budget(wb, tf32):
    """Rollout inference (no autograd): nothing is saved, every layer re-uses one set of buffers, same values."""
    from wheatley_b200.batch import jssp_batch_from_observation
    from wheatley_b200.gnn_dgl import GnnDGL
    from wheatley_b200.workload import jssp_observation_batch
    n_j, n_m, B, L = 10, 4, 2, 4
    obs = jssp_observation_batch(n_j, n_m, B, seed0=1, progress=0.2)
    torch.manual_seed(0)
    gnn = GnnDGL(obs["features"].shape[2], "gatv2", "learn", False, n_j * n_m, n_m, 3, L, 64, "gelu", 4,
                 conflicts="clique", split_output=True).cuda()
    jb, _ = jssp_batch_from_observation(obs, "cuda", n_j * n_m, conflicts="clique", graph_pooling="learn")
    nf, ge = gnn(jb)
    with torch.no_grad():
        nf2, ge2 = gnn(jb)
    assert torch.equal(nf, nf2) and torch.equal(ge, ge2)
