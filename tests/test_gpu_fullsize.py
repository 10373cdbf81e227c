"""Parity at BASELINE.json's FULL sizes through size-independent properties (the CPU oracle cannot run there):
  * additivity -- with the parameters fixed, every data-dependent sum of the path is a sum over rows, so the
    log-likelihood of the whole sample equals what the two halves give (the counter-based generator makes the
    halves bit-identical to the rows of the whole): logl(all) - const(N) = [logl(A) - const(N_A)] + [logl(B) - const(N_B)]
    reduces to logl(all) = logl(A) + logl(B) because every term of loglC_fast is linear in (N, ssq, Gram);
  * EM never decreases the log-likelihood; the loadings stay orthonormal to rounding;
  * the estimates approach the generating parameters at the 1/sqrt(N) rate (statistical check of the whole chain);
  * a fit on the first rows of the big sample equals the oracle-checked small-size path (same kernels, other grid)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SHAPES = {"c2": (1_000_000, 500, 100, 3), "c3": (10_000_000, 1000, 200, 3)}


def _free_gib():
    import torch
    return torch.cuda.mem_get_info(0)[0] / 2**30


@pytest.mark.parametrize("name", ["c2", "c3"])
def test_full_size_properties(name):
    from ppls_b200 import Data, Fit, default_context, logl_W, synth
    N, p, q, r = SHAPES[name]
    need = 8 * N * (p + q) / 2**30 * 1.02 + 2
    if _free_gib() < need:
        pytest.skip(f"{name} needs {need:.0f} GiB of HBM")
    ctx = default_context()
    truth = synth.truth_params(p, q, r)
    start = synth.start_params(truth, seed=1)
    th = (start["W"], start["C"], start["B"], start["sigE"], start["sigF"], start["sigH"], start["sigT"])
    d = Data.synthetic(ctx, N, p, q, truth, seed=20260101)
    ll_all = logl_W(None, None, *th, data=d)
    steps = 12
    f = Fit(d, start, r, steps, -np.inf, "SVD")
    f.run()
    res = f.result(want_mu=False)
    f.close()
    d.close()
    # (1) monotone, (2) orthonormal
    ll = res["loglik"]
    assert len(ll) == steps and np.all(np.diff(ll) > -1e-9 * np.abs(ll[:-1]))
    assert np.max(np.abs(res["W"].T @ res["W"] - np.eye(r))) < 1e-12
    assert np.max(np.abs(res["C"].T @ res["C"] - np.eye(r))) < 1e-12
    # (3) the chain recovers the generating parameters (sign/order free comparison through projectors)
    Pw = res["W"] @ res["W"].T; Pt = truth["W"] @ truth["W"].T
    assert np.linalg.norm(Pw - Pt) < 0.05, np.linalg.norm(Pw - Pt)
    assert abs(res["sigE"] / truth["sigE"] - 1) < 0.02 and abs(res["sigF"] / truth["sigF"] - 1) < 0.02
    # (4) additivity over the two halves of the same sample
    h = N // 2 + 12345
    parts = []
    for row0, n in ((0, h), (h, N - h)):
        dh = Data.synthetic(ctx, n, p, q, truth, seed=20260101, row0=row0)
        parts.append(logl_W(None, None, *th, data=dh))
        dh.close()
    assert abs((parts[0] + parts[1]) / ll_all - 1) < 1e-12, (parts, ll_all)


def test_prefix_of_the_big_sample_is_the_small_problem(oracle_lib):
    """First 3000 rows of the C2 sample: the device fit against the fused CPU oracle on the same rows."""
    from oracle import cbind, ppls_fused as F
    from helpers import relmax
    from ppls_b200 import Data, Fit, default_context, synth
    N, p, q, r = 3000, 500, 100, 3
    truth = synth.truth_params(p, q, r)
    st = synth.start_params(truth, seed=1)
    ctx = default_context()
    d = Data.synthetic(ctx, N, p, q, truth, seed=20260101)
    f = Fit(d, st, r, 8, -np.inf, "SVD")
    f.run()
    res = f.result(want_mu=False)
    f.close(); d.close()
    X, Y = cbind.gen_data(oracle_lib, 20260101, 0, N, truth)
    o = F.fused_loop(oracle_lib, X, Y, st["W"], st["C"], st["B"], st["sigE"], st["sigF"], st["sigH"], st["sigT"], EMsteps=8,
                     atol=-np.inf)
    assert relmax(res["loglik"], o["loglik"]) < 1e-10
    assert relmax(res["W"], o["W"]) < 1e-8 and relmax(res["C"], o["C"]) < 1e-8 and relmax(res["B"], o["B"]) < 1e-8
