"""ppls_orth_noalloc: the allocation-free variant of ppls_orth (scratch in a static device array, include/ppls_b200.h).
No counterpart in the reference -- added for callers that cannot call cudaMalloc.
Checked through the C ABI against the existing entry point ppls_orth on the same inputs, on two calls in a row (the calls
share the scratch array, which keeps the last call's contents), both orth types; sizes beyond the static maximum are
refused with a clear error.  The file name sorts last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_ELEMS = 65536          # PPLS_ORTH_NOALLOC_MAX_ELEMS


@pytest.fixture(scope="module")
def ctx():
    import torch
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    from ppls_b200 import Context
    c = Context(0)
    yield c
    c.close()


def _noalloc(ctx, G, typ):
    """G: n x r numpy.  Device buffers come from torch (the CALLER's allocations); the entry point itself allocates nothing."""
    import torch
    from ppls_b200 import _lib as L
    n, r = G.shape
    g = torch.from_numpy(np.ascontiguousarray(G)).to("cuda:0")          # row-major [n][r]: the engine's layout
    out = torch.full((n, r), float("nan"), dtype=torch.float64, device="cuda:0")
    torch.cuda.synchronize()
    L.check(L.lib().ppls_orth_noalloc(ctx._h, g.data_ptr(), n, r, L.PPLS_ORTH_SVD if typ == "SVD" else L.PPLS_ORTH_QR,
                                      out.data_ptr()))
    return out.cpu().numpy()


@pytest.mark.parametrize("typ", ["QR", "SVD"])
def test_noalloc_orth_matches_ppls_orth_on_two_calls_in_a_row(ctx, typ):
    from ppls_b200 import orth
    rng = np.random.default_rng(11)
    # two different problems back to back: the second call finds the first one's data in the shared scratch
    for n, r in [(8001, 5), (1203, 3), (MAX_ELEMS // 8, 8)]:
        G = rng.standard_normal((n, r)) * np.geomspace(1.0, 0.05, r) * 1e2
        want = orth(G, typ, ctx=ctx)                                      # existing entry point (host in / host out, cudaMalloc inside)
        got1 = _noalloc(ctx, G, typ)
        got2 = _noalloc(ctx, G, typ)
        assert np.max(np.abs(got1 - want)) < 1e-8 and np.max(np.abs(got2 - want)) < 1e-8
        assert np.array_equal(got1, got2)                                 # same kernel, same inputs: the same bits
        assert np.max(np.abs(got1.T @ got1 - np.eye(r))) < 1e-12


def test_noalloc_orth_rejects_sizes_beyond_the_static_maximum(ctx):
    import torch
    from ppls_b200 import PPLSError, _lib as L
    n, r = MAX_ELEMS // 4 + 1, 4
    g = torch.zeros((n, r), dtype=torch.float64, device="cuda:0")
    out = torch.zeros_like(g)
    with pytest.raises(PPLSError) as ei:
        L.check(L.lib().ppls_orth_noalloc(ctx._h, g.data_ptr(), n, r, L.PPLS_ORTH_QR, out.data_ptr()))
    assert ei.value.code == L.PPLS_E_UNSUPPORTED and str(MAX_ELEMS) in str(ei.value)
    with pytest.raises(PPLSError):
        L.check(L.lib().ppls_orth_noalloc(ctx._h, None, 10, 2, L.PPLS_ORTH_QR, out.data_ptr()))
    with pytest.raises(PPLSError):
        L.check(L.lib().ppls_orth_noalloc(ctx._h, g.data_ptr(), 3, 4, L.PPLS_ORTH_QR, out.data_ptr()))   # more columns than rows
