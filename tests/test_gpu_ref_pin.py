"""GPU parity against the REFERENCE'S OWN COMPILED CODE (oracle/_ref/libppls_ref.so = Package/PPLS/src/loglC.cpp
compiled unmodified, see tests/test_ref_pin.py), through the C ABI:
  ppls_loglik                      vs loglC_fast                      (loglC.cpp:319-338)  1e-10 relative
  ppls_sequential_run, every step  vs EMstepC_fast + loglC_fast       (:341-397)           1e-8 / 1e-10
  ppls_meta_run, every step        vs meta_Estep / meta_Mstep         (:400-474)           1e-8 / 1e-10
  ppls_simult_run, loglik trace    vs loglC_fast at every iterate                           1e-10
The R-level glue around those native calls (coefficients, deflation, pooling, break rules) comes from the oracle's
restatement; the native arithmetic is the reference's.  The library is prebuilt in the build container and travels
to the GPU box (it cannot be rebuilt there: /root/reference is absent)."""
import numpy as np
import pytest

from helpers import TOL_LOGLIK, TOL_PARAM, golden_cases, init_of, load_golden, loglik_tol, relmax
from oracle import ppls_meta, ppls_ref as R, ppls_seq, refbind
from test_oracle_seq import load_seq, seq_cases

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not refbind.available(), reason="oracle/_ref was not shipped")]


@pytest.fixture(scope="module")
def ctx():
    import torch
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    from ppls_b200 import Context
    c = Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("name", golden_cases())
def test_ppls_loglik_vs_reference_loglc_fast(ctx, name):
    from ppls_b200 import logl_W
    g = load_golden(name)
    for W, Cm, b, t, sg in ((g["init_W"], g["init_C"], g["init_B"], g["init_sigT"], g["init_sig"]),
                            (g["loop_W"], g["loop_C"], g["loop_B"], g["loop_sigT"], g["est_sig"])):
        co = R.coefficients_logl(b, t, sg[0], sg[1], sg[2])
        ref = refbind.loglC_fast(W, Cm, g["X"], g["Y"], sg[0], sg[1], t**2, co["c1"], co["c2"], co["c3"], co["Kc"])
        got = logl_W(g["X"], g["Y"], W, Cm, b, sg[0], sg[1], sg[2], t, ctx=ctx)
        assert abs(got / ref - 1) < loglik_tol(name)


@pytest.mark.parametrize("name", ["c1_break_rule", "r5_odd_shapes", "r1_p_lt_q", "r8_max", "r3_wide_row"])
def test_simult_trace_vs_reference_loglc_fast(ctx, name):
    """Every entry of the GPU's log-likelihood trace against the reference's loglC_fast evaluated at the oracle's iterate
    of that step (the iterates themselves are held to the oracle elsewhere)."""
    from ppls_b200 import PPLS_simult
    g = load_golden(name)
    trace = []

    def rec(i, s):
        co = R.coefficients_logl(np.diag(s["B"]), np.diag(s["sigT"]), s["sigE"], s["sigF"], s["sigH"])
        trace.append(refbind.loglC_fast(s["W"], s["C"], g["X"], g["Y"], s["sigE"], s["sigF"], np.diag(s["sigT"])**2,
                                        co["c1"], co["c2"], co["c3"], co["Kc"]))

    sg = g["init_sig"]
    R.ppls_simult_loop(g["X"], g["Y"], g["init_W"], g["init_C"], g["init_B"], sg[0], sg[1], sg[2], g["init_sigT"],
                       EMsteps=int(g["EMsteps"]), atol=float(g["atol"]), type=str(g["type"]), record=rec)
    out = PPLS_simult(g["X"], g["Y"], int(g["r"]), EMsteps=int(g["EMsteps"]), atol=float(g["atol"]), type=str(g["type"]),
                      init=init_of(g), ctx=ctx, want_mu=False)
    assert out["loglik"].shape == (len(trace),)
    assert relmax(out["loglik"], np.array(trace)) < TOL_LOGLIK


@pytest.mark.parametrize("name", seq_cases())
def test_sequential_fit_vs_reference_emstepc_fast(ctx, name, monkeypatch):
    """ppls_sequential_run against PPLS (EM_W_multi.R:229-279) driven by the reference's EMstepC_fast and loglC_fast."""
    from ppls_b200 import PPLS
    g = load_seq(name)
    monkeypatch.setattr(ppls_seq, "emstepc_fast", refbind.EMstepC_fast)
    monkeypatch.setattr(R, "loglc_fast", refbind.loglC_fast)
    ref = ppls_seq.ppls(g["X"], g["Y"], int(g["a"]), int(g["EMsteps"]), float(g["atol"]), g["guess"], g["custom"])
    out = PPLS(g["X"], g["Y"], int(g["a"]), int(g["EMsteps"]), float(g["atol"]), g["guess"], g["custom"], ctx=ctx,
               loglikelihoods=False)
    assert np.array_equal(out["Other_output"]["Number_steps"], ref["Number_steps"])
    assert relmax(out["W"], ref["W"]) < TOL_PARAM and relmax(out["C"], ref["C"]) < TOL_PARAM
    assert relmax(out["B"], ref["B"]) < TOL_PARAM and np.max(np.abs(out["sig"] / ref["sig"] - 1)) < TOL_PARAM
    for i, t in enumerate(out["Other_output"]["logvalues"]):
        assert relmax(t, ref["logvalues"][i][:t.size]) < TOL_LOGLIK


def test_one_sequential_step_vs_reference(ctx):
    """ONE EM step of one component from a custom start: nothing but logl_W's coefficients (R) and EMstepC_fast."""
    from ppls_b200 import PPLS
    g = load_seq("seq_custom_2")
    X, Y = g["X"], g["Y"]
    rng = np.random.default_rng(4)
    w = rng.standard_normal(X.shape[1]); c = rng.standard_normal(Y.shape[1])
    guess = dict(W=w / np.linalg.norm(w), C=c / np.linalg.norm(c), B=0.9, sigE=0.6, sigF=0.8, sigH=0.4, sigT=1.3)
    co = R.coefficients_estep([guess["B"]], [guess["sigT"]], guess["sigE"], guess["sigF"], guess["sigH"])
    ref = refbind.EMstepC_fast(guess["W"], guess["C"], guess["B"], X, Y, guess["sigE"], guess["sigF"], guess["sigH"],
                               guess["sigT"], co["c1"][0], co["c2"][0], co["c3"][0])
    out = PPLS(X, Y, 1, 1, -np.inf, "custom", guess, ctx=ctx, loglikelihoods=False)
    assert relmax(out["W"][:, 0], ref["W"]) < TOL_PARAM and relmax(out["C"][:, 0], ref["C"]) < TOL_PARAM
    assert abs(out["B"][0] / ref["B"] - 1) < TOL_PARAM
    want = np.array([ref["sighat"][0], ref["sighat"][1], ref["siglathat"][0], ref["siglathat"][1]])
    assert np.max(np.abs(out["sig"][0] / want - 1)) < TOL_PARAM


def test_meta_fit_vs_reference_meta_steps(ctx, monkeypatch):
    from ppls_b200 import meta_PPLSi
    from test_ref_pin import _meta_problem
    X, Y, counts = _meta_problem()
    monkeypatch.setattr(ppls_meta, "meta_estep", lambda *a: dict(refbind.meta_step(*a), _args=a))
    monkeypatch.setattr(ppls_meta, "meta_mstep", lambda e: refbind.meta_step(*e["_args"], mstep=True))
    monkeypatch.setattr(R, "loglc_fast", refbind.loglC_fast)
    for steps in (1, 12):
        ref = ppls_meta.meta_pplsi(X, Y, counts, EMsteps=steps, atol=-np.inf)
        out = meta_PPLSi(X, Y, np.repeat(np.arange(len(counts)), counts), steps, -np.inf, ctx=ctx)
        assert out["Number_steps"] == ref["Number_steps"] == steps
        assert relmax(out["W"], ref["W"]) < TOL_PARAM and relmax(out["C"], ref["C"]) < TOL_PARAM
        for a, b in zip(out["params"], ref["params"]):
            for key in ("B_T", "sigX", "sigY", "sigH", "sigT"):
                assert abs(a[key] / b[key] - 1) < TOL_PARAM, key
        assert relmax(out["log"], ref["log"]) < TOL_LOGLIK
