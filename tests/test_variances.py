"""variances.PPLS_simult (EM_W_multi.R:830-860): CPU checks of the oracle restatement and GPU parity of
ppls_variances / ppls_b200.variances against it.  PARITY UNPINNED (the reference holds no fixtures)."""
import numpy as np
import pytest

from helpers import init_of, load_golden, relmax
from oracle import ppls_ref as R, ppls_var as V


def _fit_of(name, steps=40):
    g = load_golden(name)
    th = init_of(g)
    W0, C0, B0, T0 = R.normalise(th["W"], th["C"], th["B"], th["sigT"])
    fit = R.ppls_simult_loop(g["X"], g["Y"], W0, C0, B0, th["sigE"], th["sigF"], th["sigH"], T0, EMsteps=steps, atol=-np.inf)
    return g, fit


def test_oracle_matches_the_literal_n_by_n_form():
    """:846 written literally, t(X) %*% diag(Ctt, N) %*% X with the N x N matrix materialised."""
    g, fit = _fit_of("c1_wellcond_1000", 10)
    X = g["X"][:200]
    f2 = dict(Expectations={k: (v[:200] if k in ("mu_T", "mu_U") else v) for k, v in fit["Expectations"].items()},
              estimates=fit["estimates"])
    o = V.variances_ppls_simult(f2, X, "X")
    N = X.shape[0]
    Ctt = N * f2["Expectations"]["Ctt"][0, 0]
    lit = X.T @ np.diag(np.full(N, Ctt)) @ X
    assert relmax(lit, X.T @ (Ctt * X)) < 1e-13
    c = o["components"][0]
    assert np.allclose(c["SSt_exp"], c["SSt_exp"].T, rtol=0, atol=1e-12 * np.abs(c["SSt_exp"]).max())
    assert np.all(np.isfinite(o["seLoad"])) and np.all(o["seLoad"] > 0)
    assert o["seLoad"].shape == (X.shape[1], 2)
    # varMatrix is minus the inverse of B_exp - SSt_exp
    A = c["B_exp"] - c["SSt_exp"]
    assert relmax(-o["varMatrix"][0] @ A, np.eye(A.shape[0])) < 1e-10


def _check(out, ref, r):
    assert relmax(out["seLoad"], ref["seLoad"]) < 1e-8
    assert relmax(np.abs(out["W"]), np.abs(ref["W"])) < 1e-8
    for i in range(r):
        c = ref["components"][i]
        assert relmax(out[i]["SSt_exp"], c["SSt_exp"]) < 1e-10
        assert relmax(out[i]["SSt_star"], c["SSt_star"]) < 1e-9
        assert relmax(out[i]["B_exp"], c["B_exp"]) < 1e-12
        assert relmax(out["varMatrix"][i], ref["varMatrix"][i]) < 1e-8


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["c1_wellcond_1000", "r5_odd_shapes", "r1_p_lt_q", "r8_max"])
@pytest.mark.parametrize("side", ["X", "Y"])
def test_gpu_variances_golden_inputs(name, side):
    from ppls_b200 import variances
    g, fit = _fit_of(name, 15)
    dat = g["X"] if side == "X" else g["Y"]
    ref = V.variances_ppls_simult(fit, dat, side)
    out = variances(fit, dat, side)
    _check(out, ref, ref["seLoad"].shape[1])


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(3000, 300, 150, 3), (1500, 129, 257, 2)])
@pytest.mark.parametrize("panel", ["0", "1"])
def test_gpu_variances_several_tiles(shape, panel, monkeypatch):
    """More than one 128-column block (off-diagonal Gram tiles), row-major and panel-major storage."""
    import os, sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
    from make_golden_seq import model_data
    from ppls_b200 import variances
    monkeypatch.setenv("PPLS_PANEL", panel)
    n, p, q, r = shape
    X, Y = model_data(5 + n, n, p, q, r, 0.5)
    g = np.random.default_rng(3)
    W0 = np.linalg.qr(g.standard_normal((p, r)))[0]; C0 = np.linalg.qr(g.standard_normal((q, r)))[0]
    fit = R.ppls_simult_loop(X, Y, W0, C0, np.ones(r), 0.6, 0.6, 0.3, np.ones(r) * 1.5, EMsteps=6, atol=-np.inf)
    for side, dat in (("X", X), ("Y", Y)):
        ref = V.variances_ppls_simult(fit, dat, side)
        out = variances(fit, dat, side)
        _check(out, ref, r)


@pytest.mark.gpu
def test_gpu_variances_on_a_gpu_fit_and_resident_data():
    """The whole chain on the device: PPLS_simult -> variances with the resident Data handle."""
    from ppls_b200 import Data, PPLS_simult, default_context, variances
    g = load_golden("c1_wellcond_1000")
    ctx = default_context()
    d = Data.from_host(ctx, g["X"], g["Y"])
    fit = PPLS_simult(None, None, 2, EMsteps=30, atol=-np.inf, init=init_of(g), data=d)
    out = variances(fit, None, "X", dev_data=d, matrices=False)
    ref = V.variances_ppls_simult(fit, g["X"], "X")
    d.close()
    assert set(out) == {"seLoad", "W"} and relmax(out["seLoad"], ref["seLoad"]) < 1e-8
    with pytest.raises(ValueError):
        variances(fit, g["X"], "Z")
