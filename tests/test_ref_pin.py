"""Pins the CPU oracle -- and with it every golden vector -- to the REFERENCE'S OWN COMPILED CODE.

oracle/_ref/libppls_ref.so is /root/reference/Package/PPLS/src/loglC.cpp compiled unmodified (recipe oracle/Makefile
`ref`, stand-in Eigen/Rcpp headers in oracle/ref_shim/: those libraries are not installed here).  These tests hold
  * oracle/ppls_ref.py::loglc_fast           to loglC_fast      (loglC.cpp:319-338)   SURVEY 8a row a6
  * oracle/ppls_seq.py::emstepc_fast         to EMstepC_fast    (loglC.cpp:341-397)   SURVEY 8f row 1
  * oracle/ppls_meta.py::meta_estep / mstep  to meta_Estep / meta_Mstep (:400-474)    SURVEY 8f row 4
  * the closed-form log-likelihood logl_W    to the dense Cholesky form of loglC (:53-88)           rows a5 + a6
  * the fast E-step moments of Expect_M      to the dense E-step EMstepC (:117-250)                 row a2
and replay whole golden trajectories (PPLS_simult log-likelihood traces, sequential and meta fits) with the
reference's functions swapped in for ours.  Thresholds: 1e-13 relative where the same formula is evaluated (the
stand-in's matrix product sums in index order, Eigen's and numpy's in blocks: a few ulp), 1e-9 where a dense
O((p+q)^3) form is compared with a closed form.  What stays unpinned (R code, no interpreter): the R-level
expression order of Expect_M / Maximiz_M (a2, a3), the normalisation (a9) and OmicsPLS::orth (a4).
"""
import math

import numpy as np
import pytest

from helpers import golden_cases, load_golden, relmax
from oracle import ppls_meta, ppls_ref as R, ppls_seq, refbind
from test_oracle_seq import load_seq, seq_cases

pytestmark = pytest.mark.skipif(not refbind.available(), reason="oracle/_ref not built and /root/reference absent")


def _rel(a, b):
    return abs(a - b) / max(abs(b), 1e-300)


def test_ref_library_is_the_reference_source():
    L = refbind.load()
    assert b"loglC.cpp (unmodified)" in L.ppls_ref_source()


@pytest.mark.parametrize("name", golden_cases())
def test_loglc_fast_matches_reference(name):
    g = load_golden(name)
    X, Y = g["X"], g["Y"]
    for W, Cm, b, t, sg in ((g["init_W"], g["init_C"], g["init_B"], g["init_sigT"], g["init_sig"]),
                            (g["loop_W"], g["loop_C"], g["loop_B"], g["loop_sigT"], g["est_sig"])):
        co = R.coefficients_logl(b, t, sg[0], sg[1], sg[2])
        ours = R.loglc_fast(W, Cm, X, Y, sg[0], sg[1], t**2, co["c1"], co["c2"], co["c3"], co["Kc"])
        ref = refbind.loglC_fast(W, Cm, X, Y, sg[0], sg[1], t**2, co["c1"], co["c2"], co["c3"], co["Kc"])
        assert _rel(ours, ref) < 1e-13


@pytest.mark.parametrize("name", golden_cases())
def test_golden_loglik_is_what_the_reference_computes(name):
    """The last entry of the committed trace, recomputed by the reference's loglC_fast at the committed final state."""
    g = load_golden(name)
    sg, t = g["est_sig"], g["loop_sigT"]
    co = R.coefficients_logl(g["loop_B"], t, sg[0], sg[1], sg[2])
    ref = refbind.loglC_fast(g["loop_W"], g["loop_C"], g["X"], g["Y"], sg[0], sg[1], t**2, co["c1"], co["c2"], co["c3"], co["Kc"])
    assert _rel(float(g["loglik"][-1]), ref) < 1e-12


@pytest.mark.parametrize("name", ["c1_break_rule", "r5_odd_shapes", "r3_qr", "r1_p_lt_q", "one_step"])
def test_simult_trace_with_reference_loglik(name, monkeypatch):
    """PPLS_simult's loop with the reference's loglC_fast in place of ours reproduces the golden trace and stops at
    the same step (the break rule reads these values, EM_W_multi.R:792)."""
    g = load_golden(name)
    monkeypatch.setattr(R, "loglc_fast", refbind.loglC_fast)
    sg = g["init_sig"]
    out = R.ppls_simult_loop(g["X"], g["Y"], g["init_W"], g["init_C"], g["init_B"], sg[0], sg[1], sg[2], g["init_sigT"],
                             EMsteps=int(g["EMsteps"]), atol=float(g["atol"]), type=str(g["type"]))
    assert out["loglik"].shape == g["loglik"].shape
    assert relmax(out["loglik"], g["loglik"]) < 1e-12


def _coef1(B, sigX, sigY, sigH, sigT):
    co = R.coefficients_estep([B], [sigT], sigX, sigY, sigH)
    return float(co["c1"][0]), float(co["c2"][0]), float(co["c3"][0])


@pytest.mark.parametrize("name", seq_cases())
def test_emstepc_fast_matches_reference(name):
    g = load_seq(name)
    X, Y = g["X"], g["Y"]
    W, Cm, B = g["W"][:, 0], g["C"][:, 0], float(g["B"][0])
    sigX, sigY, sigH, sigT = (float(v) for v in g["sig"][0])
    c1, c2, c3 = _coef1(B, sigX, sigY, sigH, sigT)
    ours = ppls_seq.emstepc_fast(W, Cm, B, X, Y, sigX, sigY, sigH, sigT, c1, c2, c3)
    ref = refbind.EMstepC_fast(W, Cm, B, X, Y, sigX, sigY, sigH, sigT, c1, c2, c3)
    for k in ("B", "Cut", "Ctt", "Cuu", "Cee", "Cff", "Chh"):
        assert _rel(ours[k], ref[k]) < 1e-12, k
    for k in ("mu_T", "mu_U", "W", "C", "sighat", "siglathat"):
        assert relmax(ours[k], ref[k]) < 1e-12, k


@pytest.mark.parametrize("name", seq_cases())
def test_sequential_golden_with_reference_step(name, monkeypatch):
    """The whole sequential fit (PPLS, EM_W_multi.R:229-279) with the reference's EMstepC_fast AND loglC_fast swapped
    in reproduces the committed golden vectors: same step counts, parameters, log-likelihood traces."""
    g = load_seq(name)
    monkeypatch.setattr(ppls_seq, "emstepc_fast", refbind.EMstepC_fast)
    monkeypatch.setattr(R, "loglc_fast", refbind.loglC_fast)
    fit = ppls_seq.ppls(g["X"], g["Y"], int(g["a"]), int(g["EMsteps"]), float(g["atol"]), g["guess"], g["custom"])
    assert np.array_equal(fit["Number_steps"], g["Number_steps"])
    assert relmax(fit["W"], g["W"]) < 1e-9 and relmax(fit["C"], g["C"]) < 1e-9
    assert relmax(fit["B"], g["B"]) < 1e-9 and relmax(fit["sig"], g["sig"]) < 1e-9
    for i, t in enumerate(fit["logvalues"]):
        assert relmax(t, g["logvalues"][i, :t.size]) < 1e-11


def _meta_problem(seed=5, counts=(150, 90, 120), p=14, q=9):
    rng = np.random.default_rng(seed)
    N = sum(counts)
    w = rng.standard_normal(p); w /= np.linalg.norm(w)
    c = rng.standard_normal(q); c /= np.linalg.norm(c)
    T = rng.standard_normal(N) * np.repeat([1.0, 1.4, 0.8], counts)
    U = T * np.repeat([1.2, 0.7, 1.0], counts) + 0.3 * rng.standard_normal(N)
    X = np.outer(T, w) + 0.4 * rng.standard_normal((N, p))
    Y = np.outer(U, c) + 0.5 * rng.standard_normal((N, q))
    return X, Y, list(counts)


def test_meta_estep_mstep_match_reference():
    X, Y, counts = _meta_problem()
    p, q = X.shape[1], Y.shape[1]
    W, Cm, B, sg, sl = ppls_seq.initial_guess(p, q, "equal")
    c1, c2, c3 = _coef1(B, sg[0], sg[1], sl[0], sl[1])
    Xi, Yi = X[:counts[0]], Y[:counts[0]]
    ours = ppls_meta.meta_estep(W, Cm, B, Xi, Yi, sg[0], sg[1], sl[0], sl[1], c1, c2, c3)
    ref = refbind.meta_step(W, Cm, B, Xi, Yi, sg[0], sg[1], sl[0], sl[1], c1, c2, c3)
    for k in ("Cut", "Ctt", "Cee", "Cff", "Chh"):
        assert _rel(ours[k], ref[k]) < 1e-12, k
    for k in ("Cxt", "Cyu", "mu_T", "mu_U"):
        assert relmax(ours[k], ref[k]) < 1e-12, k
    om = ppls_meta.meta_mstep(ours)
    rm = refbind.meta_step(W, Cm, B, Xi, Yi, sg[0], sg[1], sl[0], sl[1], c1, c2, c3, mstep=True)
    assert _rel(om["B"], rm["B"]) < 1e-12
    assert relmax(om["sighat"], rm["sighat"]) < 1e-12 and relmax(om["siglathat"], rm["siglathat"]) < 1e-12
    assert relmax(om["Cxt"], rm["Cxt"]) < 1e-12 and relmax(om["Cyu"], rm["Cyu"]) < 1e-12


def test_meta_fit_with_reference_steps(monkeypatch):
    """meta_PPLSi (EM_W_multi.R:509-589) run twice: our restatement, and with the reference's meta_Estep, meta_Mstep
    and loglC_fast swapped in."""
    X, Y, counts = _meta_problem()
    ours = ppls_meta.meta_pplsi(X, Y, counts, EMsteps=15, atol=-np.inf)
    monkeypatch.setattr(ppls_meta, "meta_estep",
                        lambda *a: dict(refbind.meta_step(*a), _args=a))
    monkeypatch.setattr(ppls_meta, "meta_mstep", lambda e: refbind.meta_step(*e["_args"], mstep=True))
    monkeypatch.setattr(R, "loglc_fast", refbind.loglC_fast)
    ref = ppls_meta.meta_pplsi(X, Y, counts, EMsteps=15, atol=-np.inf)
    assert ours["Number_steps"] == ref["Number_steps"] == 15
    assert relmax(ours["W"], ref["W"]) < 1e-10 and relmax(ours["C"], ref["C"]) < 1e-10
    assert relmax(ours["log"], ref["log"]) < 1e-12
    for a, b in zip(ours["params"], ref["params"]):
        for k in a:
            assert _rel(a[k], b[k]) < 1e-10, k


def _zero_orth(p, q):
    """Orthogonal parts with zero variance: the PO2PLS model of loglC / EMstepC then IS the PPLS model (their
    orthogonal terms vanish identically).  One Y-orthogonal column; q X-orthogonal columns (P_Xosc = I_q), because
    EMstepC writes `sig2Uo * P_Xosc` (loglC.cpp:201, ny x ny times q x ny), which only has matching inner dimensions
    when ny == q."""
    e1 = np.zeros((p, 1)); e1[0, 0] = 1.0
    return e1, np.eye(q), np.zeros((1, 1)), np.zeros((q, q))


@pytest.mark.parametrize("name", ["c1_wellcond_1000", "r5_odd_shapes", "r1_p_lt_q", "one_step", "r8_max"])
def test_closed_form_loglik_vs_reference_dense_cholesky(name):
    """logl_W + loglC_fast (closed-form determinant and inverse) against the reference's own dense evaluation
    (loglC.cpp:53-88: full covariance, LLT, row-wise quadratic forms).  The dense branch's constant is
    -0.5*N*p*log(2 pi) (:81, X side only); loglC_fast's is -0.5*N*(p+q)*log(2 pi) (:336)."""
    g = load_golden(name)
    X, Y = g["X"], g["Y"]
    N, p = X.shape; q = Y.shape[1]
    # one component: the dense branch multiplies C (q x a) by the 1 x 1 sig2H (loglC.cpp:43,59), so the reference's
    # own code only runs for a = 1 -- any column of the golden state is a valid one-component parameter set
    W, Cm, b, t, sg = g["loop_W"][:, :1], g["loop_C"][:, :1], g["loop_B"][:1], g["loop_sigT"][:1], g["est_sig"]
    PY, PX, so, su = _zero_orth(p, q)
    dense = refbind.loglC(W, Cm, PY, PX, b, np.hstack([X, Y]), sg[0], sg[1], sg[2], t, so, su)["value"]
    dense += -0.5 * N * q * math.log(2 * math.pi)
    fast = R.logl_w(X, Y, W, Cm, b, sg[0], sg[1], sg[2], t)
    # W, C of the golden state are orthonormal to rounding, which the closed form assumes
    assert _rel(fast, dense) < 1e-9


def _square_problem(seed, N, p, r, sig):
    """p == q: EMstepC forms Cff from Identity(p,p) (loglC.cpp:221), so it only runs for square shapes."""
    from ppls_b200 import synth
    rng = np.random.default_rng(seed)
    truth = synth.truth_params(p, p, r, *sig)
    T = rng.standard_normal((N, r)) * truth["sigT"]
    U = T * truth["B"] + truth["sigH"] * rng.standard_normal((N, r))
    X = T @ truth["W"].T + truth["sigE"] * rng.standard_normal((N, p))
    Y = U @ truth["C"].T + truth["sigF"] * rng.standard_normal((N, p))
    return X, Y, synth.start_params(truth, seed)


@pytest.mark.parametrize("seed,N,p,r,sig", [(1, 120, 12, 1, (0.5, 0.5, 0.3)), (2, 90, 9, 1, (0.3, 0.4, 0.2)),
                                            (3, 60, 6, 1, (0.05, 0.05, 0.05))])
def test_fast_estep_moments_vs_reference_dense_estep(seed, N, p, r, sig):
    """Expect_M's fast branch (EM_W_multi.R:668-716, oracle expect_m) against the reference's compiled dense E-step
    EMstepC (loglC.cpp:117-250; one component and p == q: it multiplies C by a 1 x 1 sig2H, :123,:143, and builds Cff
    from Identity(p,p), :221) at an orthonormal start: Ctt, Cuu, Cut, Chh, tr(Cee)/p, tr(Cff)/q and
    X'mu_T / N, Y'mu_U / N (the inputs of Maximiz_M's orth, :732-733)."""
    X, Y, st = _square_problem(seed, N, p, r, sig)
    W, Cm, b, t = st["W"], st["C"], np.asarray(st["B"], float), np.asarray(st["sigT"], float)
    fast = R.expect_m(X, Y, W, Cm, b, st["sigE"], st["sigF"], st["sigH"], t)
    PY, PX, so, su = _zero_orth(p, p)
    dense = refbind.EMstepC(W, Cm, PY, PX, b, np.hstack([X, Y]), st["sigE"], st["sigF"], st["sigH"], t, so, su)
    assert relmax(np.diag(fast["Ctt"]), np.abs(np.diag(dense["Ctt"]))) < 1e-9
    assert relmax(np.diag(fast["Cuu"]), np.abs(np.diag(dense["Cuu"]))) < 1e-9
    assert relmax(np.diag(fast["Cut"]), np.diag(dense["Cut"])) < 1e-9
    assert relmax(fast["Chh"], np.abs(dense["Chh"])) < 1e-9
    assert _rel(fast["Cee"][0, 0], float(np.mean(dense["diag_Cee"]))) < 1e-9
    assert _rel(fast["Cff"][0, 0], float(np.mean(dense["diag_Cff"]))) < 1e-9
    assert relmax(X.T @ fast["mu_T"] / N, dense["Cxt"]) < 1e-9
    assert relmax(Y.T @ fast["mu_U"] / N, dense["Cyu"]) < 1e-9


def test_simulc_model_structure():
    """simulC (loglC.cpp:269-315) compiled from the reference, on the stand-in rnorm: the sample covariance matches the
    model covariance of sseXY_W plus the orthogonal parts -- the same check tests/test_simulc.py makes on OUR generator
    (R's RNG stream itself cannot be reproduced outside R, so only the model is pinned).  One component: the
    reference's `VectorXd U = Tt * B_T + H` (:309) only type-checks for a = 1."""
    p, q, N = 6, 5, 200000
    rng = np.random.default_rng(0)
    W = np.linalg.qr(rng.standard_normal((p, 1)))[0]; Cm = np.linalg.qr(rng.standard_normal((q, 1)))[0]
    PY = np.linalg.qr(rng.standard_normal((p, 1)))[0]; PX = np.linalg.qr(rng.standard_normal((q, 1)))[0]
    B, sigX, sigY, sigH, sigT, sTo, sUo = 1.3, 0.4, 0.5, 0.3, 1.1, 0.7, 0.6
    D = refbind.simulC(N, W, Cm, PY, PX, [B], sigX, sigY, sigH, [sigT], [sTo], [sUo], seed=3)
    S = R.sse_xy_w(W, Cm, [B], sigX, sigY, sigH, [sigT])
    S[:p, :p] += sTo**2 * PY @ PY.T
    S[p:, p:] += sUo**2 * PX @ PX.T
    emp = D.T @ D / N
    assert np.max(np.abs(emp - S)) < 0.02
    with pytest.raises(RuntimeError):          # two components: the reference's own code rejects its VectorXd
        W2 = np.linalg.qr(rng.standard_normal((p, 2)))[0]; C2 = np.linalg.qr(rng.standard_normal((q, 2)))[0]
        refbind.simulC(10, W2, C2, PY, PX, [1.0, 0.5], sigX, sigY, sigH, [1.0, 0.8], [sTo], [sUo])
