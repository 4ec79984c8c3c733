"""``b200gat_colsum_noalloc`` (include/b200gat.h): the column-sum kernel pair with its slab partials in a file-scope
device array instead of a caller-provided scratch buffer, for callers that cannot allocate.  Checked through the C-ABI
against ``b200gat_colsum`` on two calls in a row (the second with a different shape, so it reuses the array while it
still holds the first call's partials).  Collected last on purpose (the file name sorts after every other test file)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _reference(_lib, x, M, N, ld):
    out = torch.full((N,), float("nan"), dtype=torch.float32, device=x.device)
    scratch = torch.empty(max(1, int(_lib.lib.b200gat_colsum_scratch_elems(M, N))), dtype=torch.float32,
                          device=x.device)
    _lib.check(_lib.lib.b200gat_colsum(x.data_ptr(), M, N, ld, out.data_ptr(), scratch.data_ptr(),
                                       _lib.stream_ptr()), "b200gat_colsum")
    return out


def _noalloc(_lib, x, M, N, ld, out):
    _lib.check(_lib.lib.b200gat_colsum_noalloc(x.data_ptr(), M, N, ld, out.data_ptr(), _lib.stream_ptr()),
               "b200gat_colsum_noalloc")


def test_colsum_noalloc_matches_colsum_on_two_calls_in_a_row():
    from wheatley_b200 import _lib
    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cpu").manual_seed(7)
    cases = [(16008, 256, 256), (777, 100, 104)]          # (M, N, ld): the bench's bias gradient, then a ragged one
    xs = [torch.randn(M, ld, generator=g).to(dev) for M, _, ld in cases]
    outs = [torch.full((N,), float("nan"), dtype=torch.float32, device=dev) for _, N, _ in cases]
    refs = [_reference(_lib, x, M, N, ld) for x, (M, N, ld) in zip(xs, cases)]
    torch.cuda.synchronize()
    # inputs and outputs exist: the calls under test must not allocate device memory
    before = torch.cuda.memory_stats(dev)["num_device_alloc"] if hasattr(torch.cuda, "memory_stats") else None
    free0, _ = torch.cuda.mem_get_info(dev)
    for x, out, (M, N, ld) in zip(xs, outs, cases):        # two calls in a row, same stream, no sync in between
        _noalloc(_lib, x, M, N, ld, out)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(dev)
    assert free1 >= free0, (free0, free1)
    if before is not None:
        assert torch.cuda.memory_stats(dev)["num_device_alloc"] == before
    for out, ref, x, (M, N, ld) in zip(outs, refs, xs, cases):
        assert torch.equal(out, ref)                       # same slab decomposition, same summation order: bitwise
        exact = x[:, :N].double().sum(0)
        err = (out.double() - exact).abs().max().item() / max(exact.abs().max().item(), 1e-30)
        assert err <= 1e-4, err                            # the fp32 tolerance of the parity tests
    # and once more over the first shape: the array now holds the second call's partials
    again = torch.full_like(outs[0], float("nan"))
    _noalloc(_lib, xs[0], *cases[0], again)
    torch.cuda.synchronize()
    assert torch.equal(again, refs[0])


def test_colsum_noalloc_rejects_widths_beyond_the_static_scratch_and_handles_empty_input():
    from wheatley_b200 import _lib
    dev = torch.device("cuda", 0)
    max_n = 1024                                            # B200GAT_COLSUM_NOALLOC_MAX_N (include/b200gat.h)
    x = torch.ones(8, max_n + 4, device=dev)
    out = torch.zeros(max_n + 4, device=dev)
    rc = _lib.lib.b200gat_colsum_noalloc(x.data_ptr(), 8, max_n + 4, max_n + 4, out.data_ptr(), _lib.stream_ptr())
    assert rc != 0
    msg = _lib.lib.b200gat_last_error().decode()
    assert "B200GAT_COLSUM_NOALLOC_MAX_N" in msg and str(max_n + 4) in msg, msg
    with pytest.raises(_lib.B200GatError):
        _lib.check(rc, "b200gat_colsum_noalloc")
    # the maximum width itself is accepted
    out_ok = torch.full((max_n,), float("nan"), device=dev)
    _noalloc(_lib, x, 8, max_n, max_n + 4, out_ok)
    # M = 0: zeros, no kernel
    out0 = torch.full((16,), float("nan"), device=dev)
    _noalloc(_lib, x, 0, 16, 16, out0)
    torch.cuda.synchronize()
    assert torch.equal(out_ok, torch.full((max_n,), 8.0, device=dev))
    assert torch.equal(out0, torch.zeros(16, device=dev))
    assert C.sizeof(C.c_float) * 2 * 148 * max_n <= 2 << 20   # the documented, modest size of the static array
