"""Boundary behaviours SURVEY 8b lists: the fused layers re-executed under
``torch.utils.checkpoint`` (psp/models/gnn_flat.py:192-202), wrapped by ``torch.compile(dynamic=True)``
(psp/models/gnn_mp.py:266), and an in-place ``features += residual`` on the layer's output (gnn_flat.py:230)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def wb():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import wheatley_b200
    return wheatley_b200


def _graph(N=300, E=2000, seed=0):
    g = torch.Generator().manual_seed(seed)
    src, dst = torch.randint(0, N, (E,), generator=g), torch.randint(0, N, (E,), generator=g)
    ei = torch.stack([torch.cat([src, torch.arange(N)]), torch.cat([dst, torch.arange(N)])]).cuda()
    return ei, torch.randn(N, 64, generator=g).cuda(), torch.randn(ei.shape[1], 64, generator=g).cuda()


def test_graphconv_under_activation_checkpointing(wb):
    from torch.utils.checkpoint import checkpoint
    ei, x, ea = _graph()
    torch.manual_seed(0)
    layer = wb.GraphConv(64, 64, 4, pyg=True, edge_dim=64).cuda()
    xa = x.clone().requires_grad_(True)
    out_a, _ = layer.forward_nog(xa, ei, ea)
    out_a.square().sum().backward()
    ga = {k: p.grad.clone() for k, p in layer.named_parameters()}
    for p in layer.parameters():
        p.grad = None
    xb = x.clone().requires_grad_(True)
    out_b = checkpoint(lambda t: layer.forward_nog(t, ei, ea)[0], xb, use_reentrant=False)
    out_b.square().sum().backward()
    assert torch.equal(out_a, out_b)
    assert torch.allclose(xa.grad, xb.grad, rtol=1e-5, atol=1e-6)
    for k, p in layer.named_parameters():
        assert torch.allclose(p.grad, ga[k], rtol=2e-4, atol=1e-5), k      # atomically accumulated ones: order noise


def test_graphconv_wrapped_by_torch_compile(wb):
    """``torch.compile(module, dynamic=True)`` as the reference wraps its GNN: the fused ops are opaque to the
    tracer (``torch.compiler.disable``) and run eagerly inside the wrapper; same values, gradients flow."""
    ei, x, ea = _graph(seed=1)
    torch.manual_seed(0)
    layer = wb.GraphConv(64, 64, 4, pyg=True, edge_dim=64).cuda()
    ref, _ = layer.forward_nog(x, ei, ea)
    compiled = torch.compile(layer, dynamic=True)
    xc = x.clone().requires_grad_(True)
    out, _ = compiled.forward_nog(xc, ei, ea) if hasattr(compiled, "forward_nog") else compiled(None, xc, ea)
    assert torch.equal(out, ref)
    out.sum().backward()
    assert xc.grad is not None and torch.isfinite(xc.grad).all()
    mlp = wb.MLP(3, 256, 256, 64, False, "gelu").cuda()
    h = torch.randn(500, 256, device="cuda")
    assert torch.allclose(torch.compile(mlp, dynamic=True)(h), mlp(h), rtol=1e-5, atol=1e-5)


def test_inplace_residual_add_on_the_layer_output(wb):
    """gnn_flat.py:230 adds the residual IN PLACE into the layer output: the output must not alias what the backward
    reads (``agg_save`` is a private copy)."""
    ei, x, ea = _graph(seed=2)
    torch.manual_seed(0)
    layer = wb.GraphConv(64, 64, 4, pyg=True, edge_dim=64).cuda()
    R = torch.randn(x.shape[0], 256, device="cuda")
    xa = x.clone().requires_grad_(True)
    (layer.forward_nog(xa, ei, ea)[0] + R).square().sum().backward()
    xb = x.clone().requires_grad_(True)
    out, _ = layer.forward_nog(xb, ei, ea)
    out += R                                             # in place, like the reference
    out.square().sum().backward()
    assert torch.allclose(xa.grad, xb.grad, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("flavour", ["jssp", "psp"])
def test_batches_are_freed_by_reference_counting(flavour):
    """No reference cycle through a batch, its plan and the plan's clique view: dropping the last reference frees the
    device tensors at once.  (A plan <-> view cycle and ``PspGraphData._graph = self`` used to leave every minibatch
    to the cyclic garbage collector: the caching allocator grew by ~3 MiB per step and stalled for 10-200 ms whenever a
    late collection returned the lot.)"""
    import gc
    import weakref
    if flavour == "jssp":
        from wheatley_b200.batch import jssp_batch_from_observation
        from wheatley_b200.workload import jssp_observation_batch
        obs = jssp_observation_batch(10, 4, 3, seed0=0, progress=0.3)

        def build():
            jb, _ = jssp_batch_from_observation(obs, "cuda", 40, conflicts="clique", graph_pooling="learn")
            return jb, jb.plan
    else:
        from wheatley_b200.psp_batch import preprocess
        from wheatley_b200.wire import PspArena
        from wheatley_b200.workload import psp_taillard_graph
        arena = PspArena.from_graphs([psp_taillard_graph(4, 4, b, progress=0.0) for b in range(8)])

        def build():
            tup = preprocess(arena, "cuda", conflicts="clique", graph_pooling="learn")
            return tup, tup[0].plan
    build()                                   # warm caches (library handles, allocator)
    gc.collect()
    gc.disable()
    try:
        holder, plan = build()
        view = plan.clique_view(plan.sort_edge_ints(torch.zeros(plan.n_edges, dtype=torch.int64, device="cuda")))
        refs = [weakref.ref(plan), weakref.ref(view), weakref.ref(plan.in_ptr)]
        assert view.plan is plan
        del holder, plan, view
        assert all(r() is None for r in refs), [type(r()).__name__ for r in refs if r() is not None]
    finally:
        gc.enable()
