"""CPU-only: the C-ABI library loads and exports every symbol include/b200gat.h declares;
host-side logic (edge-feature factorisation, sharding, flat gradient bucket over gloo)."""
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built():
    import __graft_entry__ as ge
    ge.build()
    import wheatley_b200
    return wheatley_b200


def test_library_exports_every_declared_symbol(built):
    import ctypes
    from wheatley_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "b200gat.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(b200gat_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in include/b200gat.h but not exported"
    assert declared == set(_lib.SYMBOLS), declared ^ set(_lib.SYMBOLS)
    assert _lib.lib.b200gat_version() >= 100
    assert _lib.launch_count() == 0          # nothing may launch without a GPU


def test_argument_validation_without_gpu(built):
    from wheatley_b200 import _lib
    rc = _lib.lib.b200gat_plan_build(None, None, -1, 3, None, None, None, None, None, None, None, None, None)
    assert rc == 1 and b"negative" in _lib.lib.b200gat_last_error()
    rc = _lib.lib.b200gat_mask_head_fwd(None, 0, None, 0, 2, 3, -1e20, None, None, None, None)
    assert rc == 1


def test_no_cpu_fallback(built):
    wb = built
    with pytest.raises(wb.B200GatError):
        wb.GraphPlan(torch.tensor([0]), torch.tensor([0]), 1)
    with pytest.raises(wb.B200GatError):
        wb.masked_log_softmax(torch.randn(2, 3))


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "wheatley_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), f
                assert "/root/reference" not in txt, f


def test_psp_edge_embedder_factorisation_matches_reference_golden(built, golden_dir):
    from wheatley_b200.edge_feats import PspEdgeEmbedder
    g = np.load(os.path.join(golden_dir, "edge_embedder.npz"))
    H = g["out"].shape[1]
    emb = PspEdgeEmbedder("sum", 4, H, "all", False)
    emb.load_state_dict({k[2:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("p.")})
    ef = emb.factor(torch.from_numpy(g["type"]), torch.from_numpy(g["att_rc"]), torch.from_numpy(g["att_rp"]))
    np.testing.assert_allclose(ef.dense().detach().numpy(), g["out"], rtol=1e-6, atol=1e-6)
    assert ef.attrs.shape[1] == 5 and ef.attr_weight.shape == (H, 5)


def test_mlp_matches_reference_golden(built, golden_dir):
    g = np.load(os.path.join(golden_dir, "mlp.npz"))
    n = len({k.split(".")[0] for k in g.files})
    for i in range(n):
        act, norm = str(g[f"{i}.act"]), bool(g[f"{i}.norm"])
        sd = {k[len(f"{i}.p."):]: torch.from_numpy(g[k]) for k in g.files if k.startswith(f"{i}.p.")}
        m = built.MLP(3, 10, 12, 5, norm, act)
        m.load_state_dict(sd)
        x = torch.from_numpy(g[f"{i}.x"]).requires_grad_(True)
        y = m(x)
        np.testing.assert_allclose(y.detach().numpy(), g[f"{i}.y"], rtol=1e-5, atol=1e-6)
        (y * torch.from_numpy(g[f"{i}.R"])).sum().backward()
        np.testing.assert_allclose(x.grad.numpy(), g[f"{i}.gx"], rtol=1e-4, atol=1e-6)


def test_state_dict_names_match_reference(built):
    wb = built
    assert sorted(wb.EEGATConv(4, 4, 3, 3, 2, edge_scoring=True).state_dict()) == sorted(
        ["fc_node_src.weight", "fc_ni.weight", "fc_nj.weight", "fc_edge_src.weight", "fc_fij.weight",
         "attn", "fc_fij_e.weight", "attn_e", "bias"])          # generic/eegatconv.py:134-162
    assert sorted(wb.GATv3Conv(4, 3, 2, edge_dim=4).state_dict()) == sorted(
        ["context_attention_layer.weight", "context_attention_layer.bias", "lin_l.weight", "lin_l.bias",
         "lin_r.weight", "lin_r.bias", "lin_edge.weight"])
    from wheatley_b200.gnn_dgl import GnnDGL
    keys = set(GnnDGL(10, "gatv2", "learn", False, 20, 4, 3, 2, 8, "gelu", 2, conflicts="clique").state_dict())
    for k in ("node_embedder.weight", "edge_embedder.0.weight", "embedder.layers.0.weight",
              "features_extractors.1.fc_fij.weight", "features_extractors.0.attn", "mlps.1.layers.2.bias"):
        assert k in keys, k


def test_workload_sizes_match_survey(built):
    from wheatley_b200.workload import jssp_batch
    for (nj, nm), (n, e) in {(10, 10): (101, 1281), (15, 15): (226, 4021), (30, 20): (601, 19741)}.items():
        b = jssp_batch(nj, nm, 1)
        assert (b["n_nodes"], b["n_edges"]) == (n, e)


def test_workload_batch_matches_oracle_batching(built):
    from oracle import batch_oracle
    from wheatley_b200.workload import jssp_batch, jssp_observation
    obs = [jssp_observation(6, 4, s, progress=0.5) for s in range(3)]
    maxe = max(o["n_edges"] for o in obs)
    feat = np.stack([o["features"] for o in obs])
    ei = np.zeros((3, 2, maxe), np.int64)
    for i, o in enumerate(obs):
        ei[i, :, : o["n_edges"]] = o["edge_index"]
    ob = batch_oracle.jssp_batch(feat, ei, [o["n_nodes"] for o in obs], [o["n_edges"] for o in obs],
                                 conflicts="clique")
    ob = batch_oracle.jssp_add_pool_and_machine_nodes(ob, "learn", "clique")
    wbat = jssp_batch(6, 4, 3, progress=0.5)
    assert np.array_equal(ob["src"], wbat["src"].numpy())
    assert np.array_equal(ob["dst"], wbat["dst"].numpy())
    assert np.array_equal(ob["type"], wbat["etype"].numpy())


def test_batch_sizes_from_observation_match_oracle(built):
    """Host half of the device batch builder: per-graph node / edge offsets and clique edge counts computed from
    the observation alone (no device work) equal what the oracle's builder produces."""
    import torch
    from oracle import batch_oracle
    from wheatley_b200.batch import jssp_sizes
    rng = np.random.default_rng(5)
    B, n, F = 4, 23, 9
    nn = np.array([23, 1, 17, 9])
    ne = np.array([5, 0, 30, 2])
    feat = rng.standard_normal((B, n, F)).astype(np.float32)
    mid = rng.integers(0, 4, (B, n)).astype(np.float32)
    mid[rng.random((B, n)) < 0.3] = -1
    feat[:, :, 6] = mid
    ei = np.zeros((B, 2, 30), np.int64)
    for conflicts in ("att", "clique"):
        ob = batch_oracle.jssp_batch(feat, ei, nn, ne, conflicts=conflicts)
        meta, n_mid = jssp_sizes(torch.from_numpy(feat), nn, ne, conflicts)
        assert meta[0].tolist() == [0] + np.cumsum(nn).tolist()
        assert meta[1].tolist() == [0] + np.cumsum(ob["batch_num_edges"]).tolist()
        assert meta[2, :B].tolist() == ne.tolist()
        if conflicts == "clique":
            assert n_mid == int(mid[np.arange(n)[None, :] < nn[:, None]].max()) + 1
            assert meta[3, :B].tolist() == (ob["batch_num_edges"] - 2 * ne - nn).tolist()


def test_sharding(built):
    from wheatley_b200.ppo import shard_by_edges, shard_range
    for n, w in [(10, 4), (3, 8), (128, 8), (7, 1)]:
        r = [shard_range(n, k, w) for k in range(w)]
        assert r[0][0] == 0 and r[-1][1] == n and all(a[1] == b[0] for a, b in zip(r, r[1:]))
    rng = np.random.default_rng(0)
    for w in (1, 2, 4, 8):
        e = rng.integers(100, 5000, 64).tolist()
        sh = shard_by_edges(e, w)
        assert sh[0][0] == 0 and sh[-1][1] == 64 and all(a[1] == b[0] for a, b in zip(sh, sh[1:]))
        assert all(hi > lo for lo, hi in sh)
        loads = [sum(e[lo:hi]) for lo, hi in sh]
        assert max(loads) <= sum(e) / w + max(e)


_GLOO = r'''
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, %r)
rank, world = int(sys.argv[1]), int(sys.argv[2])
os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=sys.argv[3], RANK=str(rank), WORLD_SIZE=str(world))
dist.init_process_group("gloo", rank=rank, world_size=world)
import wheatley_b200
from wheatley_b200.ppo import FlatGradAllReduce, shard_range
torch.manual_seed(0)
model = wheatley_b200.MLP(3, 6, 8, 2, False, "gelu")
bucket = FlatGradAllReduce(model.parameters())
x = torch.randn(10, 6)
lo, hi = shard_range(10, rank, world)
bucket.zero()
(model(x[lo:hi]).sum() / 10 * world).backward()        # each rank: its shard of the global mean
bucket.reduce()
ref = wheatley_b200.MLP(3, 6, 8, 2, False, "gelu")
ref.load_state_dict(model.state_dict())
(ref(x).sum() / 10).backward()
for p, q in zip(model.parameters(), ref.parameters()):
    assert p.grad.data_ptr() >= bucket.flat.data_ptr()
    assert torch.allclose(p.grad, q.grad, atol=1e-6), (p.grad - q.grad).abs().max()
# FlatAdam on the all-reduced bucket == torch.optim.Adam on the global-batch gradients, for 3 steps
from wheatley_b200.ppo import FlatAdam
ref_opt = torch.optim.Adam(ref.parameters(), lr=1e-2)
opt = FlatAdam(model.parameters(), lr=1e-2, bucket=bucket)
for it in range(3):
    if it:
        opt.zero_grad()
        (model(x[lo:hi]).sum() / 10 * world).backward()
        bucket.reduce()
        ref_opt.zero_grad()
        (ref(x).sum() / 10).backward()
    opt.step()
    ref_opt.step()
    for p, q in zip(model.parameters(), ref.parameters()):
        assert torch.allclose(p, q, atol=1e-6), (it, (p - q).abs().max())
dist.destroy_process_group()
print("ok", rank)
'''


def test_flat_gradient_allreduce_world_size_2_gloo(built, tmp_path):
    script = tmp_path / "gloo_worker.py"
    script.write_text(_GLOO % ROOT)
    port = str(29500 + os.getpid() % 2000)
    procs = [subprocess.Popen([sys.executable, str(script), str(r), "2", port], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT) for r in range(2)]
    outs = [p.communicate(timeout=240)[0].decode() for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o


def test_flat_adam_matches_torch_adam(built):
    """ppo.FlatAdam (one flat parameter, one concatenated gradient) against torch.optim.Adam parameter by
    parameter, including a parameter that never receives a gradient and the state dict round trip."""
    import torch
    import wheatley_b200
    from wheatley_b200.ppo import FlatAdam
    torch.manual_seed(0)
    a = wheatley_b200.MLP(3, 6, 8, 2, False, "gelu")
    a.unused = torch.nn.Parameter(torch.randn(5))
    b = wheatley_b200.MLP(3, 6, 8, 2, False, "gelu")
    b.unused = torch.nn.Parameter(a.unused.detach().clone())
    b.load_state_dict(a.state_dict())
    oa = FlatAdam(a.parameters(), lr=3e-3, weight_decay=0.0)
    ob = torch.optim.Adam(b.parameters(), lr=3e-3)
    assert all(p.data_ptr() >= oa.flat.data_ptr() for p in a.parameters())      # parameters live in the flat buffer
    x = torch.randn(12, 6)
    for it in range(4):
        oa.zero_grad()
        ob.zero_grad()
        a(x).square().mean().backward()
        b(x).square().mean().backward()
        oa.step()
        ob.step()
        for (k, p), q in zip(a.named_parameters(), b.parameters()):
            assert torch.allclose(p, q, atol=1e-7, rtol=1e-6), (it, k)
    assert torch.equal(a.unused, b.unused)
    sd = oa.state_dict()
    oa.load_state_dict(sd)
    assert set(a.state_dict()) == set(b.state_dict())


def test_gemm_descriptor_validation_without_gpu(built):
    """b200gat_gemm_supported / workspace sizing are host-only: shapes the tcgen05 GEMM covers, the ones it
    hands to the library GEMM, and the split-K workspace of a weight-gradient GEMM (include/b200gat.h)."""
    from wheatley_b200 import _lib
    P = 1 << 20                      # fake, 16-byte aligned device addresses (never dereferenced on the host)

    def desc(a_major, b_major, M, N, K, lda, ldb, ldd, **kw):
        g = _lib.Gemm()
        g.a_major, g.b_major, g.M, g.N, g.K = a_major, b_major, M, N, K
        g.A, g.lda, g.B, g.ldb, g.D, g.ldd = P, lda, 2 * P, ldb, 3 * P, ldd
        g.a_parts, g.d_parts = kw.get("a_parts", 1), kw.get("d_parts", 1)
        g.a_part_stride, g.d_part_stride = kw.get("a_part_stride", 0), kw.get("d_part_stride", 0)
        g.epilogue = kw.get("epilogue", _lib.EPI_NONE)
        g.D2, g.Zin, g.ld2 = kw.get("D2"), kw.get("Zin"), kw.get("ld2", 0)
        return g

    sup = _lib.lib.b200gat_gemm_supported
    assert sup(desc(0, 0, 16008, 256, 256, 256, 256, 256)) == 1            # MLP layer forward
    assert sup(desc(0, 0, 16000, 64, 464, 464, 464, 64)) == 1              # head layer 1 (K = 464)
    assert sup(desc(0, 1, 16000, 464, 64, 64, 464, 464)) == 1              # its dx: ragged 464-wide output
    assert sup(desc(0, 0, 16008, 768, 64, 64, 64, 256, d_parts=3, d_part_stride=16008 * 256)) == 1
    assert sup(desc(0, 0, 10, 1, 64, 64, 64, 1)) == 0                      # width-1 head output
    assert sup(desc(0, 0, 10, 64, 30, 30, 30, 64)) == 0                    # 30-wide rows: not 16-byte aligned
    assert sup(desc(0, 0, 10, 64, 64, 64, 64, 64, epilogue=_lib.EPI_GELU)) == 0       # GELU needs D2
    assert sup(desc(0, 0, 10, 64, 64, 64, 64, 64, epilogue=_lib.EPI_GELU, D2=4 * P, ld2=64)) == 1
    g = desc(0, 0, 10, 64, 64, 64, 64, 64)
    g.A = P + 4                                                             # misaligned operand
    assert sup(g) == 0
    # weight gradient dW[256, 256] = dy^T x over 16,008 rows: split-K, partials + row sums in the workspace
    g = desc(1, 1, 256, 256, 16008, 256, 256, 256)
    ws = _lib.lib.b200gat_gemm_workspace_elems(g)
    assert ws > 0 and ws % (256 * 256) == 0 and ws // (256 * 256) >= 16
    g.a_rowsum = 5 * P
    assert _lib.lib.b200gat_gemm_workspace_elems(g) == ws // (256 * 256) * (256 * 256 + 256)
    assert _lib.lib.b200gat_gemm_workspace_elems(desc(0, 0, 16008, 256, 256, 256, 256, 256)) == 0
    assert _lib.launch_count() == 0


def test_batched_edge_tables_match_per_layer_linear(built):
    """edge_feats.edge_tables (one GEMM for all layers' projected edge-type tables) against the per-layer
    ``fc_fij(Embedding.weight) + bias`` it replaces: values and every gradient (pure torch: runs on the CPU)."""
    import torch
    from wheatley_b200.edge_feats import edge_tables
    torch.manual_seed(0)
    T, F, HD, L = 6, 16, 24, 3
    for with_bias in (True, False):
        emb = torch.randn(T, F, dtype=torch.float64, requires_grad=True)
        Ws = [torch.randn(HD, F, dtype=torch.float64, requires_grad=True) for _ in range(L)]
        bs = [torch.randn(HD, dtype=torch.float64, requires_grad=True) if with_bias else None for _ in range(L)]
        tabs = edge_tables(emb, Ws, bs)
        coef = [torch.randn(T, HD, dtype=torch.float64) for _ in range(L)]
        # layer 1's table is left unused: its gradient arrives as None
        loss = sum((tabs[id(Ws[l])] * coef[l]).sum() for l in (0, 2))
        loss.backward()
        got = [emb.grad.clone()] + [w.grad for w in Ws] + [b.grad if with_bias else None for b in bs]
        emb.grad = None
        for w in Ws:
            w.grad = None
        for b in bs:
            if b is not None:
                b.grad = None
        ref_tabs = [torch.nn.functional.linear(emb, Ws[l], bs[l]) for l in range(L)]
        for l in range(L):
            assert torch.allclose(tabs[id(Ws[l])], ref_tabs[l], rtol=1e-12, atol=1e-12)
        sum((ref_tabs[l] * coef[l]).sum() for l in (0, 2)).backward()
        ref = [emb.grad] + [w.grad for w in Ws] + [b.grad if with_bias else None for b in bs]
        for g, r in zip(got, ref):
            if r is None:
                assert g is None or float(g.abs().max()) == 0.0
            else:
                assert torch.allclose(g, r, rtol=1e-12, atol=1e-12)
    assert edge_tables(emb, [], []) is None
    assert edge_tables(emb, Ws, [bs[0], None, None]) is None or not with_bias


def test_split_rows_backward_is_a_concatenation(built):
    import torch
    from wheatley_b200.gnn_dgl import _SplitRows
    x = torch.randn(7, 3, dtype=torch.float64, requires_grad=True)
    a, b = _SplitRows.apply(x, 5)
    assert torch.equal(a, x[:5]) and torch.equal(b, x[5:])
    (a * 2).sum().backward()
    assert torch.equal(x.grad, torch.cat([torch.full((5, 3), 2.0, dtype=torch.float64), torch.zeros(2, 3, dtype=torch.float64)]))
    x.grad = None
    a, b = _SplitRows.apply(x, 5)
    ((a * 2).sum() + (b * 3).sum()).backward()
    assert torch.equal(x.grad[:5], torch.full((5, 3), 2.0, dtype=torch.float64)) and float(x.grad[5:].min()) == 3.0


def test_lazy_categorical_is_the_reference_distribution(built):
    import torch
    from torch.distributions.categorical import Categorical
    from wheatley_b200.agent_head import LazyCategorical
    logits = torch.randn(4, 9, requires_grad=True)
    lazy, ref = LazyCategorical(logits), Categorical(logits=logits)
    assert lazy._d is None
    assert torch.equal(lazy.probs, ref.probs) and torch.equal(lazy.logits, ref.logits)
    assert torch.equal(lazy.entropy(), ref.entropy())
    a = torch.tensor([0, 3, 8, 2])
    assert torch.equal(lazy.log_prob(a), ref.log_prob(a))
    lazy.probs.sum().backward()                  # stays in the autograd graph of the logits
    assert logits.grad is not None
    with __import__("pytest").raises(AttributeError):
        lazy._missing


def test_batch_builder_refuses_cpu_and_unknown_modes(built):
    """The device batch builder has no CPU path and says so; option validation happens before any device work."""
    import torch
    import wheatley_b200 as wb
    from wheatley_b200.batch import jssp_batch_on_device
    from wheatley_b200.prefetch import BatchPrefetcher
    obs = dict(features=torch.zeros(1, 2, 8), edge_index=torch.zeros(1, 2, 1, dtype=torch.int64),
               n_nodes=torch.tensor([2]), n_edges=torch.tensor([1]))
    with pytest.raises(wb.B200GatError):
        jssp_batch_on_device(obs, "cpu", conflicts="clique")
    with pytest.raises(wb.B200GatError):
        BatchPrefetcher("cpu")
    if torch.cuda.is_available():
        return
    # without a GPU the mode checks still run first when the device is a CUDA device name
    with pytest.raises(ValueError):
        jssp_batch_on_device(obs, "cuda:0", conflicts="node")            # needs max_n_machines
    with pytest.raises(ValueError):
        jssp_batch_on_device(obs, "cuda:0", conflicts="bogus")


def test_wire_format_round_trip_and_gather(built):
    """Observation wire format: encode -> decode is the identity on every field, slots are
    independent, capacities are enforced, and a minibatch gather is a plain copy of whole slots."""
    import numpy as np
    from wheatley_b200.wire import PspArena, PspObsLayout
    from wheatley_b200.workload import psp_taillard_graph
    graphs = [psp_taillard_graph(4, 4, 1, n_rp=3), psp_taillard_graph(3, 2, 2), psp_taillard_graph(5, 3, 3, n_rp=1)]
    layout = PspObsLayout.for_graphs(graphs)
    assert layout.slot_words % 4 == 0
    arena = PspArena(layout, 5, pin=False)
    for i, g in enumerate(graphs):
        arena.write(i + 1, g)
    for i, g in enumerate(graphs):
        back = arena.read(i + 1)
        for k in ("feat", "resources", "res_cal", "prec", "rp", "rp_r"):
            assert np.array_equal(back[k], g[k]), k
    arena.write(1, graphs[1])                                   # overwrite with a smaller graph: no stale rows
    back = arena.read(1)
    assert back["feat"].shape == graphs[1]["feat"].shape and np.array_equal(back["prec"], graphs[1]["prec"])
    assert not arena.words[1, layout.off_res: layout.off_cal].view(np.float32).reshape(
        layout.max_nodes, layout.max_res)[graphs[1]["feat"].shape[0]:].any()
    mb = arena.gather([3, 2, 3])
    assert np.array_equal(mb.words[0], arena.words[3]) and np.array_equal(mb.words[1], arena.words[2])
    with pytest.raises(ValueError):
        PspObsLayout(4, 2, 4, 0).encode(np.zeros(1000, np.int32), **{k: graphs[0][k] for k in
                                                                  ("feat", "resources", "res_cal", "prec")})
    with pytest.raises(ValueError):
        layout.decode(np.zeros(layout.slot_words, np.int32))
    # a shared-memory buffer (what a worker process maps) works as the backing store
    import multiprocessing as mp
    shm = mp.RawArray("i", 2 * layout.slot_words)
    shared = PspArena(layout, 2, buffer=shm)
    shared.write(0, graphs[0])
    assert np.array_equal(np.frombuffer(shm, dtype=np.int32, count=layout.slot_words), shared.words[0])


def test_psp_sizes_table_matches_a_direct_count(built):
    """``psp_batch.psp_sizes`` (the host-side offset table the device builder is launched with: no device->host read)
    against a direct per-graph count and the numpy oracle's edge totals, ragged graphs, both conflict modes."""
    import numpy as np
    from oracle import psp_batch_oracle as pbo
    from wheatley_b200.psp_batch import psp_sizes
    from wheatley_b200.wire import PspArena
    from wheatley_b200.workload import psp_taillard_graph
    graphs = [psp_taillard_graph(4, 4, 1, n_rp=3), psp_taillard_graph(3, 2, 2), psp_taillard_graph(5, 3, 3, n_rp=1),
              psp_taillard_graph(6, 4, 4, progress=0.5)]
    arena = PspArena.from_graphs(graphs, pin=False)
    users = [(g["resources"] != 0).sum(0) for g in graphs]                         # per graph: users per resource
    for conflicts in ("clique", "node", "att"):
        meta, n_res, unused = psp_sizes(arena, len(graphs), conflicts)
        m = meta.numpy()
        assert m.shape == (5, len(graphs) + 1) and (m[:, 0] == 0).all()
        assert np.array_equal(np.diff(m[0]), [g["feat"].shape[0] for g in graphs])
        assert np.array_equal(np.diff(m[1]), [g["prec"].shape[1] for g in graphs])
        assert np.array_equal(np.diff(m[2]), [g["rp"].shape[1] for g in graphs])
        want_rc = [int((c * (c - 1)).sum()) if conflicts == "clique" else 0 for c in users]
        want_us = [int(c.sum()) if conflicts == "node" else 0 for c in users]
        assert np.array_equal(np.diff(m[3]), want_rc) and np.array_equal(np.diff(m[4]), want_us)
        assert n_res == max(g["resources"].shape[1] for g in graphs)
        total = np.zeros(n_res, dtype=np.int64)
        for c in users:
            total[: len(c)] += c
        assert unused == sum(1 << r for r in range(n_res) if total[r] == 0)
    ref = pbo.psp_batch(graphs, "clique", "learn", True)
    rc = int((np.asarray(ref["type"]) == 3).sum())                                 # edge type 3 = resource conflict
    assert rc == int(psp_sizes(arena, len(graphs), "clique")[0][3, -1])


def test_psp_sizes_and_oracle_agree_on_survey_table(built):
    """PyG-flavour node / edge counts per graph at t = 0 from the numpy oracle on generated graphs."""
    from oracle import psp_batch_oracle as pbo
    from wheatley_b200.workload import psp_sizes_expected, psp_taillard_graph
    for (n_j, n_m), (n_node, e_node) in {(10, 10): (113, 635), (15, 15): (243, 1400), (30, 20): (623, 3685),
                                         (4, 4): (23, 113)}.items():
        out = pbo.psp_batch([psp_taillard_graph(n_j, n_m, 0)], "node", "learn", True)
        assert (out["n_total"], out["edge_index"].shape[1]) == (n_node, e_node) == psp_sizes_expected(n_j, n_m)
    out = pbo.psp_batch([psp_taillard_graph(10, 10, 0)], "clique", "learn", True)
    assert out["n_total"] == 103 and out["edge_index"].shape[1] == psp_sizes_expected(10, 10, "clique")[1]


def test_gnn_mp_state_dict_names_match_reference_golden(built, golden_dir):
    """State-dict keys of wheatley_b200.GnnMP == the reference GnnMP's (gnn_mp.npz, incl. torch.compile's
    ``_orig_mod.`` infix, psp/models/gnn_mp.py:139,145,183,266,300), shapes included."""
    import numpy as np
    import os
    from wheatley_b200.gnn_mp import GnnMP
    g = np.load(os.path.join(golden_dir, "gnn_mp.npz"))
    for tag in ("node272", "clique272"):
        conflicts, pooling = (str(x) for x in g[f"{tag}.opts"][:2])
        Hd, H, L, max_n_nodes, max_res = (int(v) for v in g[f"{tag}.cfg"])
        gnn = GnnMP(16, pooling, max_n_nodes, max_res, 3, L, Hd, "gelu", H, conflicts=conflicts, rwpe_k=0,
                    shared_conv=False, pyg=True, graphobs=True, add_rp_edges="frontier_strict")
        sd = gnn.state_dict()
        assert sorted(sd) == sorted(str(k) for k in g[f"{tag}.state_keys"])
        for k, v in sd.items():
            assert tuple(v.shape) == g[f"{tag}.p.{k}"].shape, k
        assert gnn.features_dim == g[f"{tag}.node_features"].shape[2]


_GLOO_UNEQUAL = r'''
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, %r)
rank, world = int(sys.argv[1]), int(sys.argv[2])
os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=sys.argv[3], RANK=str(rank), WORLD_SIZE=str(world))
dist.init_process_group("gloo", rank=rank, world_size=world)
import wheatley_b200
from wheatley_b200.ppo import FlatGradAllReduce, adv_moments, shard_by_edges
torch.manual_seed(0)
model = wheatley_b200.MLP(3, 6, 8, 2, False, "gelu")
ref = wheatley_b200.MLP(3, 6, 8, 2, False, "gelu")
ref.load_state_dict(model.state_dict())
x = torch.randn(11, 6)
(ref(x).sum(1).mean()).backward()                      # the single-process global-batch mean
lo, hi = shard_by_edges([5, 1, 1, 1, 1, 1, 1, 1, 1, 1, 50], world)[rank]      # UNEQUAL shards: 10 | 1 samples
assert (lo, hi) == ((0, 10), (10, 11))[rank]
bucket = FlatGradAllReduce(model.parameters())
# (a) every rank evaluates the mean over ITS shard; the ranks' gradients are weighted by their sample counts
bucket.zero()
model(x[lo:hi]).sum(1).mean().backward()
bucket.reduce(n_local=hi - lo)
for p, q in zip(model.parameters(), ref.parameters()):
    assert torch.allclose(p.grad, q.grad, atol=1e-6), ("weighted", (p.grad - q.grad).abs().max())
# (b) every rank evaluates its SHARE of the global mean (what ppo_loss(sharded=True) returns): plain SUM
bucket.zero()
(model(x[lo:hi]).sum(1).sum() / 11).backward()
bucket.reduce(average=False)
for p, q in zip(model.parameters(), ref.parameters()):
    assert torch.allclose(p.grad, q.grad, atol=1e-6), ("sum", (p.grad - q.grad).abs().max())
# (c) the advantage statistics of the global minibatch from one all-reduce of three doubles (alg/ppo.py:602-605)
adv = torch.randn(11, generator=torch.Generator().manual_seed(3)) * 2 + 1
m = adv_moments(adv[lo:hi])
dist.all_reduce(m)
n, s1, s2 = m.tolist()
mean, std = s1 / n, ((s2 - s1 * s1 / n) / (n - 1)) ** 0.5
assert n == 11 and abs(mean - float(adv.mean())) < 1e-6 and abs(std - float(adv.std())) < 1e-6
dist.destroy_process_group()
print("ok", rank)
'''


def test_unequal_shards_world_size_2_gloo(built, tmp_path):
    """``reduce()`` divided by ``world`` whatever the shard sizes.  Unequal shards (``shard_by_edges``)
    must still give the gradient of the global-batch mean: weighted by the ranks' sample counts, or a plain SUM of
    losses that are already shares of the global mean; and the per-minibatch advantage normalisation uses the
    GLOBAL mean / std."""
    script = tmp_path / "gloo_unequal.py"
    script.write_text(_GLOO_UNEQUAL % ROOT)
    port = str(31500 + os.getpid() % 2000)
    procs = [subprocess.Popen([sys.executable, str(script), str(r), "2", port], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT) for r in range(2)]
    outs = [p.communicate(timeout=240)[0].decode() for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o


def test_flat_adam_reference_state_round_trip(built):
    """the reference resumes with ``optimizer.load_state_dict`` (alg/ppo.py:447-461).  FlatAdam exports /
    imports its moments in ``torch.optim.Adam``'s per-parameter layout."""
    import wheatley_b200
    from wheatley_b200.ppo import FlatAdam
    torch.manual_seed(0)
    a = wheatley_b200.MLP(3, 6, 8, 2, False, "gelu")
    b = wheatley_b200.MLP(3, 6, 8, 2, False, "gelu")
    b.load_state_dict(a.state_dict())
    oa, ob = FlatAdam(a.parameters(), lr=3e-3), torch.optim.Adam(b.parameters(), lr=3e-3)
    x = torch.randn(7, 6)
    for _ in range(3):
        oa.zero_grad(); ob.zero_grad()
        a(x).square().mean().backward(); b(x).square().mean().backward()
        oa.step(); ob.step()
    sd = oa.reference_state_dict()
    ref = ob.state_dict()
    for i in ref["state"]:
        assert torch.allclose(sd["state"][i]["exp_avg"], ref["state"][i]["exp_avg"], atol=1e-7)
        assert torch.allclose(sd["state"][i]["exp_avg_sq"], ref["state"][i]["exp_avg_sq"], atol=1e-9)
    # a fresh FlatAdam resumed from the reference optimizer's own state continues identically
    c = wheatley_b200.MLP(3, 6, 8, 2, False, "gelu")
    c.load_state_dict(b.state_dict())
    oc = FlatAdam(c.parameters(), lr=3e-3)
    oc.load_reference_state_dict(ref)
    oc.zero_grad(); ob.zero_grad()
    c(x).square().mean().backward(); b(x).square().mean().backward()
    oc.step(); ob.step()
    for p, q in zip(c.parameters(), b.parameters()):
        assert torch.allclose(p, q, atol=1e-6)
