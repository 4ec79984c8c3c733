"""b200gat_readout_fwd/bwd against a literal torch restatement of the reference's trunk tail:
``torch.cat`` of the per-layer tensors, per-graph ``torch.max(dim=0)`` / mean / pool-node row, ``F.pad`` to
``max_n_nodes``, ``torch.stack`` (psp/models/gnn_mp.py:419-500).  The restatement runs on the CPU, where
``torch.max`` sends the gradient to the FIRST maximal row - the behaviour the reference-generated trunk golden
(gnn_mp.npz, tag ragged_max) pins."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def wb():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import wheatley_b200
    return wheatley_b200


def reference_tail(sources, counts, max_n, pooling, pool_row0):
    """gnn_mp.py:419-470 on CPU tensors."""
    features = torch.cat(sources, dim=-1)
    num_nodes = sum(counts)
    node_features = features[:num_nodes]
    ge, nnf, start = [], [], 0
    for i, nn_ in enumerate(counts):
        rows = node_features[start:start + nn_]
        if pooling == "max":
            ge.append(torch.max(rows, dim=0)[0])
        elif pooling == "avg":
            ge.append(torch.matmul(torch.ones(nn_, dtype=rows.dtype) / nn_, rows))
        else:
            ge.append(features[pool_row0 + i])
        nnf.append(torch.nn.functional.pad(rows, (0, 0, 0, max_n - nn_)))
        start += nn_
    return torch.stack(nnf), torch.stack(ge)


@pytest.mark.parametrize("pooling", ["max", "avg", "learn"])
@pytest.mark.parametrize("counts,max_n,extra", [([5, 3, 7, 1], 8, 6), ([23] * 11, 23, 55), ([300, 17, 260], 301, 3),
                                                ([2000, 2000], 2000, 2)])
def test_readout_matches_reference_tail(wb, pooling, counts, max_n, extra):
    from wheatley_b200.readout import readout
    g = torch.Generator().manual_seed(len(counts) * 7 + max_n)
    N = sum(counts) + extra + len(counts)
    widths = [16, 64, 30, 64, 2]
    src = [torch.randn(N, w, generator=g) for w in widths]
    # ties: a column that is constant over all rows (the node-type embedding broadcast to every task node) and
    # one with repeated maxima
    src[1][:, 0] = 0.25
    src[1][:, 1] = torch.randint(0, 3, (N,), generator=g).float()
    pool_row0 = sum(counts) + extra
    ref_in = [s.clone().requires_grad_(True) for s in src]
    nf_r, ge_r = reference_tail(ref_in, counts, max_n, pooling, pool_row0)
    R1, R2 = torch.randn(nf_r.shape, generator=g), torch.randn(ge_r.shape, generator=g)
    ((nf_r * R1).sum() + (ge_r * R2).sum()).backward()
    dev_in = [s.cuda().requires_grad_(i != 0) for i, s in enumerate(src)]     # source 0: no gradient wanted
    # a strided source (column slice of a wider tensor) must work without a copy
    wide = torch.randn(N, 80, generator=g).cuda()
    wide[:, 8:72] = src[3].cuda()
    dev_in[3] = wide[:, 8:72].detach().requires_grad_(True)
    nf, ge = readout(dev_in, counts, max_n, pooling, pool_row0=pool_row0)
    assert nf.shape == nf_r.shape and ge.shape == ge_r.shape
    assert torch.equal(nf.cpu(), nf_r.detach())                                # copies / zero padding: bit-exact
    if pooling == "avg":
        assert torch.allclose(ge.cpu(), ge_r.detach(), rtol=1e-5, atol=1e-6)
    else:
        assert torch.equal(ge.cpu(), ge_r.detach())
    ((nf * R1.cuda()).sum() + (ge * R2.cuda()).sum()).backward()
    assert dev_in[0].grad is None
    for i in range(1, len(src)):
        tol = dict(rtol=1e-5, atol=1e-6) if pooling == "avg" else dict(rtol=0, atol=0)
        assert torch.allclose(dev_in[i].grad.cpu(), ref_in[i].grad, **tol), (i, pooling)


def test_readout_partial_gradients_and_errors(wb):
    from wheatley_b200.readout import readout
    x = torch.randn(10, 8, device="cuda", requires_grad=True)
    nf, ge = readout([x], [4, 5], 6, "max")
    ge.sum().backward()                      # node_features unused: its gradient never materialises
    assert x.grad.shape == (10, 8) and float(x.grad.sum()) == 16.0 and float(x.grad[9].abs().sum()) == 0.0
    nf, ge = readout([x.detach()], [4, 5], 6, None)
    assert ge is None and nf.shape == (2, 6, 8)
    with pytest.raises(wb.B200GatError):
        readout([x], [7, 2], 6, "max")       # a graph larger than max_n_nodes
    with pytest.raises(wb.B200GatError):
        readout([x.cpu()], [4, 5], 6, "max")
