"""GPU parity tests of the sequential PPLS fit (ppls_sequential_run / ppls_b200.PPLS) and of logl_W
(ppls_loglik), called through the C ABI, against oracle/ppls_seq.py and its committed golden vectors.
Tolerances: parameters 1e-8 relative, log-likelihood 1e-10 relative (BASELINE.json north_star)."""
import numpy as np
import pytest

from helpers import TOL_LOGLIK, TOL_PARAM, relmax
from test_oracle_seq import load_seq, seq_cases

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    import torch
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    from ppls_b200 import Context
    c = Context(0)
    yield c
    c.close()


def _check(out, ref_W, ref_C, ref_B, ref_sig, ref_steps, ref_logv):
    assert np.array_equal(out["Other_output"]["Number_steps"], ref_steps)
    assert relmax(out["W"], ref_W) < TOL_PARAM and relmax(out["C"], ref_C) < TOL_PARAM
    assert relmax(out["B"], ref_B) < TOL_PARAM
    assert np.max(np.abs(out["sig"] / ref_sig - 1)) < TOL_PARAM
    for i, t in enumerate(out["Other_output"]["logvalues"]):
        assert relmax(t, ref_logv[i][:t.size]) < TOL_LOGLIK


@pytest.mark.parametrize("name", seq_cases())
def test_golden_sequential(ctx, name):
    from ppls_b200 import PPLS
    g = load_seq(name)
    out = PPLS(g["X"], g["Y"], int(g["a"]), int(g["EMsteps"]), float(g["atol"]), g["guess"], g["custom"], ctx=ctx)
    assert out["class"] == "PPLS" and out["sig"].shape == (int(g["a"]), 4)
    _check(out, g["W"], g["C"], g["B"], g["sig"], g["Number_steps"], g["logvalues"])
    li = out["Other_output"]["Last_increment"]
    assert np.max(np.abs(li - g["Last_increment"])) < 1e-9 * np.max(np.abs(g["logvalues"][0, 0]))


@pytest.mark.parametrize("name", ["seq_equal_3", "seq_odd_2"])
@pytest.mark.parametrize("panel", ["0", "1"])
def test_golden_sequential_through_the_wide_path(ctx, name, panel, monkeypatch):
    from ppls_b200 import PPLS
    monkeypatch.setenv("PPLS_FORCE_WIDE", "1")
    monkeypatch.setenv("PPLS_PANEL", panel)
    g = load_seq(name)
    out = PPLS(g["X"], g["Y"], int(g["a"]), int(g["EMsteps"]), float(g["atol"]), g["guess"], g["custom"], ctx=ctx)
    _check(out, g["W"], g["C"], g["B"], g["sig"], g["Number_steps"], g["logvalues"])


@pytest.mark.parametrize("shape", [(3000, 300, 120, 3, 12), (1201, 1000, 200, 2, 8), (150, 2600, 40, 2, 6)])
def test_fresh_inputs_sequential(ctx, shape):
    """Seeded model data at sizes the numpy oracle finishes in seconds (the last one takes the wide path)."""
    import sys, os
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
    from make_golden_seq import model_data
    from oracle import ppls_seq
    from ppls_b200 import PPLS
    n, p, q, a, steps = shape
    X, Y = model_data(77 + n, n, p, q, a, 0.4)
    ref = ppls_seq.ppls(X, Y, a, steps, 1e-4, "equal")
    out = PPLS(X, Y, a, steps, 1e-4, "equal", ctx=ctx)
    logv = [t for t in ref["logvalues"]]
    _check(out, ref["W"], ref["C"], ref["B"], ref["sig"], ref["Number_steps"], logv)
    # Other_output$Loglikelihoods (:275): logl_W of the first i components on the undeflated data
    from oracle import ppls_ref
    for i in range(a):
        want = ppls_ref.logl_w(X, Y, ref["W"][:, :i + 1], ref["C"][:, :i + 1], np.diag(ref["B"][:i + 1]), ref["sig"][i, 0],
                               ref["sig"][i, 1], ref["sig"][i, 2], np.diag(ref["sig"][:i + 1, 3]))
        assert abs(out["Other_output"]["Loglikelihoods"][i] / want - 1) < 1e-9


def test_logl_w_matches_oracle(ctx):
    from oracle import ppls_ref
    from ppls_b200 import logl_W
    from helpers import init_of, load_golden
    g = load_golden("r5_odd_shapes")
    th = init_of(g)
    want = ppls_ref.logl_w(g["X"], g["Y"], th["W"], th["C"], th["B"], th["sigE"], th["sigF"], th["sigH"], th["sigT"])
    got = logl_W(g["X"], g["Y"], th["W"], th["C"], th["B"], th["sigE"], th["sigF"], th["sigH"], th["sigT"], ctx=ctx)
    assert abs(got / want - 1) < TOL_LOGLIK


def test_vanishing_noise_stops_the_fit(ctx):
    """PPLSi returns NA when a noise s.d. is below 100 eps (:153-155); PPLS keeps the earlier components
    and warns (:257-262).  Start the second component from such a state."""
    from ppls_b200 import PPLS
    g = load_seq("seq_odd_2")
    p, q = g["X"].shape[1], g["Y"].shape[1]
    ok = dict(W=np.ones(p) / np.sqrt(p), C=np.ones(q) / np.sqrt(q), B=1.0, sigE=1.0 / p, sigF=1.0 / q, sigH=1.0, sigT=1.0)
    bad = dict(ok, sigE=1e-15)
    with pytest.warns(UserWarning, match="calculations are stopped"):
        out = PPLS(g["X"], g["Y"], 2, 5, 1e-4, "custom", [ok, bad], ctx=ctx)
    assert out["W"].shape == (p, 1) and out["B"].shape == (1,) and out["sig"].shape == (1, 4)
    with pytest.warns(UserWarning):
        out0 = PPLS(g["X"], g["Y"], 1, 5, 1e-4, "custom", bad, ctx=ctx)
    assert out0["W"].shape == (p, 0)


def test_sequential_argument_errors(ctx):
    from ppls_b200 import PPLS, PPLSError
    g = load_seq("seq_p_lt_q_1")
    with pytest.raises(ValueError):
        PPLS(g["X"], g["Y"], 0, ctx=ctx)
    with pytest.raises(ValueError):
        PPLS(g["X"], g["Y"], 6, ctx=ctx)           # ncol(X) = 5 < nr_comp
    with pytest.raises(ValueError):
        PPLS(g["X"], g["Y"], 1, initialGuess="o2m", ctx=ctx)        # needs OmicsPLS::o2m: pass its result as customGuess
    with pytest.raises(PPLSError):
        PPLS(g["X"], g["Y"], 1, EMsteps=0, ctx=ctx)


def test_random_initial_guess_is_the_reference_distribution_and_the_oracle_bits(ctx):
    """initialGuess = "random" (EM_W_multi.R:133) drawn on the device: the same numbers as the host restatement
    (oracle/cbind.py::random_guess), unit-norm positive loadings, chi-square scalars with the reference's means
    (B ~ chisq_1, sigH, sigT ~ chisq_100/100, sigE, sigF ~ chisq_10/100), different for every (seed, component)."""
    import ctypes as C
    from oracle import cbind
    from ppls_b200 import _lib as L
    from ppls_b200.api import _guess
    Lo = cbind.load()
    p, q = 257, 33
    draws = []
    for seed, comp in [(7, 0), (7, 1), (8, 0)] + [(100 + s, 0) for s in range(60)]:
        g = _guess(p, q, "random", None, ctx=ctx, seed=seed, component=comp)
        o = cbind.random_guess(Lo, p, q, seed, comp)
        assert np.max(np.abs(g.W[:, 0] - o["W"])) < 1e-15 and np.max(np.abs(g.C[:, 0] - o["C"])) < 1e-15
        assert (g.B[0], g.sigT[0], g.struct.sigE, g.struct.sigF, g.struct.sigH) == (o["B"], o["sigT"], o["sigE"], o["sigF"], o["sigH"])
        assert abs(np.sum(g.W**2) - 1) < 1e-14 and np.all(g.W > 0) and abs(np.sum(g.C**2) - 1) < 1e-14
        draws.append((g.B[0], g.sigT[0], g.struct.sigH, g.struct.sigE, g.struct.sigF))
    d = np.array(draws)
    assert len({tuple(r) for r in draws}) == len(draws)
    assert 0.5 < d[:, 0].mean() < 1.6 and np.all(np.abs(d[:, 1:3].mean(axis=0) - 1.0) < 0.06)
    assert np.all(np.abs(d[:, 3:5].mean(axis=0) - 0.1) < 0.02)
    eq = _guess(p, q, "equal", None, ctx=ctx)
    assert np.allclose(eq.W, 1 / np.sqrt(p)) and eq.struct.sigE == 1.0 / p and eq.B[0] == 1.0


@pytest.mark.parametrize("name,seed", [("seq_equal_3", 1), ("seq_odd_2", 5)])
def test_sequential_fit_from_random_starts(ctx, name, seed):
    from oracle import ppls_seq
    from ppls_b200 import PPLS
    g = load_seq(name)
    a = int(g["a"])
    ref = ppls_seq.ppls(g["X"], g["Y"], a, 20, 1e-4, "random", seed=seed)
    out = PPLS(g["X"], g["Y"], a, 20, 1e-4, "random", ctx=ctx, seed=seed, loglikelihoods=False)
    _check(out, ref["W"], ref["C"], ref["B"], ref["sig"], ref["Number_steps"], ref["logvalues"])


def test_constraints_and_critfunc(ctx):
    """PPLS(..., critfunc, constraints) (EM_W_multi.R:229-230; fconstraint :85-92, applied at :142-146 and :160-167; critfunc
    at :172) against the oracle's restatement: fixed scalars per component, fixed loadings on the first component,
    critfunc = abs, and the refusal of fixed loadings that are not orthogonal to the earlier components."""
    from oracle import ppls_seq
    from ppls_b200 import PPLS, PPLSError, fconstraint
    g = load_seq("seq_equal_3")
    X, Y = g["X"], g["Y"]
    p, q = X.shape[1], Y.shape[1]
    rng = np.random.default_rng(3)
    w = rng.standard_normal(p); w /= np.linalg.norm(w)
    cases = [
        ([dict(sigH=0.2), dict(B=0.7, sigE=0.3), {}], None),
        ([dict(W=w, sigT=1.1), dict(sigF=0.25), dict(sigH=0.1, B=1.3)], "abs"),
        ([dict(C=np.ones(q) / np.sqrt(q)), {}, {}], None),
    ]
    for cons, crit in cases:
        ref = ppls_seq.ppls(X, Y, 3, 25, 1e-4, "equal", critfunc=(abs if crit else None), constraints=cons)
        out = PPLS(X, Y, 3, 25, 1e-4, "equal", ctx=ctx, loglikelihoods=False, critfunc=crit,
                   constraints=[fconstraint(c) for c in cons])
        _check(out, ref["W"], ref["C"], ref["B"], ref["sig"], ref["Number_steps"], ref["logvalues"])
    out = PPLS(X, Y, 1, 10, 1e-4, "equal", ctx=ctx, loglikelihoods=False, constraints=[dict(W=w, B=0.9, sigE=0.4)])
    assert np.array_equal(out["W"][:, 0], w) and out["B"][0] == 0.9 and out["sig"][0, 0] == 0.4
    with pytest.raises(PPLSError):                 # second component's fixed loadings not orthogonal to the first's estimate
        PPLS(X, Y, 2, 5, 1e-4, "equal", ctx=ctx, loglikelihoods=False, constraints=[{}, dict(W=w)])
    with pytest.raises(ValueError):
        PPLS(X, Y, 2, 5, ctx=ctx, constraints=[{}])                 # one list entry per component (:240)
    with pytest.raises(ValueError):
        PPLS(X, Y, 1, 5, ctx=ctx, critfunc=lambda x: x * x)


def test_ppls_simult_whole_call_with_its_own_start(ctx):
    """PPLS_simult(X, Y, a, ...) without init=: the reference's start fit PPLS(X, Y, a, 20, 1e-4, 'random') (:762-771, drawn
    from `seed` on the counter-based generator), start normalisation (:773-778), loop (:780-793), result normalisation
    (:794-799) -- against the same chain of oracle functions; also start='equal'."""
    from oracle import ppls_ref, ppls_seq
    from ppls_b200 import PPLS_simult
    g = load_seq("seq_equal_3")
    X, Y, a = g["X"], g["Y"], 3
    f0 = ppls_seq.ppls(X, Y, a, 20, 1e-4, "random", seed=11)
    W0, C0, B0, T0 = ppls_ref.normalise(f0["W"], f0["C"], f0["B"], f0["sig"][:, 3])
    ref = ppls_ref.ppls_simult_loop(X, Y, W0, C0, B0, f0["sig"][a - 1, 0], f0["sig"][a - 1, 1], f0["sig"][a - 1, 2], T0,
                                    EMsteps=30, atol=1e-4)
    out = PPLS_simult(X, Y, a, EMsteps=30, atol=1e-4, ctx=ctx, seed=11)
    # a random start is far from the fixed point: the rounding differences of the 3 x 20-step start fit are amplified over
    # the first steps of the loop (1.4e-10 at step 1-4 here) before the iteration contracts them again (1e-12 at the end)
    assert len(out["loglik"]) == len(ref["loglik"]) and relmax(out["loglik"], ref["loglik"]) < 1e-9
    assert relmax(out["loglik"][-5:], ref["loglik"][-5:]) < TOL_LOGLIK
    for k in ("W", "C", "B", "sigT"):
        assert relmax(out["estimates"][k], ref["estimates"][k]) < TOL_PARAM, k
    f0 = ppls_seq.ppls(X, Y, a, 20, 1e-4, "equal")
    W0, C0, B0, T0 = ppls_ref.normalise(f0["W"], f0["C"], f0["B"], f0["sig"][:, 3])
    ref = ppls_ref.ppls_simult_loop(X, Y, W0, C0, B0, f0["sig"][a - 1, 0], f0["sig"][a - 1, 1], f0["sig"][a - 1, 2], T0,
                                    EMsteps=30, atol=1e-4)
    out = PPLS_simult(X, Y, a, EMsteps=30, atol=1e-4, ctx=ctx, start="equal")
    assert len(out["loglik"]) == len(ref["loglik"]) and relmax(out["loglik"], ref["loglik"]) < TOL_LOGLIK
    for k in ("W", "C", "B", "sigT"):
        assert relmax(out["estimates"][k], ref["estimates"][k]) < TOL_PARAM, k
    for k in ("sigE", "sigF", "sigH"):
        assert abs(out["estimates"][k] / ref["estimates"][k] - 1) < TOL_PARAM, k
