"""CPU tests of the oracle itself (no GPU): the op-for-op restatement against independent
dense forms, against the committed golden vectors, and the fused algebra the GPU engine uses
against the op-for-op restatement (incl. virtual-rank row sharding)."""
import numpy as np
import pytest

from oracle import cbind, ppls_fused as F, ppls_ref as R
from ppls_b200 import synth
from helpers import golden_cases, init_of, load_golden, loglik_tol, moment_tol, relmax, TOL_PARAM


def _problem(oracle_lib, N, p, q, r, sig, seed=7):
    truth = synth.truth_params(p, q, r, *sig)
    X, Y = cbind.gen_data(oracle_lib, seed, 0, N, truth)
    return truth, X, Y, synth.start_params(truth, seed)


@pytest.mark.parametrize("shape", [(500, 20, 10, 2), (120, 7, 13, 1), (90, 31, 17, 5), (64, 12, 12, 3)])
def test_fast_estep_matches_dense_debug_form(oracle_lib, shape):
    """Expect_M fast branch (EM_W_multi.R:668-713) == Expect_M(debug=TRUE) (:643-667) on the
    diagonals the masks keep (:715-716), for orthonormal W, C."""
    N, p, q, r = shape
    truth, X, Y, st = _problem(oracle_lib, N, p, q, r, (0.4, 0.3, 0.2))
    args = (X, Y, st["W"], st["C"], truth["B"], 0.4, 0.3, 0.2, truth["sigT"])
    f, d = R.expect_m(*args), R.expect_m_dense(*args)
    for k in ("mu_T", "mu_U", "Ctt", "Cuu", "Cut", "Cee", "Cff"):
        assert relmax(f[k], d[k]) < 1e-10, k
    assert relmax(np.diag(f["Chh"]), np.diag(d["Chh"])) < 1e-10


@pytest.mark.parametrize("shape", [(500, 20, 10, 2), (77, 9, 14, 3)])
def test_loglik_matches_dense_gaussian(oracle_lib, shape):
    """logl_W + loglC_fast (low-rank determinant + Woodbury quadratic form) == the dense
    Gaussian log-likelihood under sseXY_W."""
    N, p, q, r = shape
    truth, X, Y, st = _problem(oracle_lib, N, p, q, r, (0.3, 0.4, 0.2))
    a = R.logl_w(X, Y, st["W"], st["C"], truth["B"], 0.3, 0.4, 0.2, truth["sigT"])
    b = R.dense_loglik(X, Y, st["W"], st["C"], truth["B"], 0.3, 0.4, 0.2, truth["sigT"])
    assert abs(a - b) / abs(b) < 1e-11


def test_closed_form_inverse_identity(oracle_lib):
    """Package/rank_one_inverse.R:56: Sigma^-1 = blkdiag(I/sE2, I/sF2) - [W c1 W', W c2 C'; C c2 W', C c3 C']."""
    truth, X, Y, st = _problem(oracle_lib, 50, 20, 10, 2, (0.3, 0.4, 0.2))
    W, C = st["W"], st["C"]
    co = R.coefficients_estep(truth["B"], truth["sigT"], 0.3, 0.4, 0.2)
    S = R.sse_xy_w(W, C, truth["B"], 0.3, 0.4, 0.2, truth["sigT"])
    c1, c2, c3 = (np.diag(co[k]) for k in ("c1", "c2", "c3"))
    inv = np.block([[np.eye(20) / 0.09 - W @ c1 @ W.T, -W @ c2 @ C.T], [-C @ c2 @ W.T, np.eye(10) / 0.16 - C @ c3 @ C.T]])
    assert np.max(np.abs(inv @ S - np.eye(30))) < 1e-10


def test_coefficients_two_spellings_agree():
    """logl_W (:312-320, g via sqrt) and Expect_M (:672-680) evaluate the same coefficients."""
    b = np.array([1.5, 1.1, 0.8]); t = np.array([1.0, 0.9, 0.8])
    a, c = R.coefficients_estep(b, t, 0.3, 0.2, 0.1), R.coefficients_logl(b, t, 0.3, 0.2, 0.1)
    for k in ("c1", "c2", "c3", "Kc"):
        assert relmax(a[k], c[k]) < 1e-13


def test_normalise_quirk():
    """:794-799 -- W, C: permute columns, then scale column k by the UNPERMUTED sign k; B: sign then permute."""
    W = np.arange(6.0).reshape(3, 2) + 1; C = np.arange(4.0).reshape(2, 2) + 1
    Wn, Cn, Bn, Tn = R.normalise(W, C, np.array([-0.5, 2.0]), np.array([1.0, 1.0]))
    assert np.allclose(Wn, np.c_[-W[:, 1], W[:, 0]])      # rot = [1, 0]; sign = [-1, +1] applied positionally
    assert np.allclose(np.diag(Bn), [2.0, 0.5]) and np.allclose(np.diag(Tn), [1.0, 1.0])
    from ppls_b200 import normalise as eng_normalise
    for a_, b_ in zip(eng_normalise(W, C, np.array([-0.5, 2.0]), np.array([1.0, 1.0])), (Wn, Cn, Bn, Tn)):
        assert np.array_equal(a_, b_)


def test_orth_qr_sign_convention():
    rng = np.random.default_rng(3)
    G = rng.standard_normal((30, 4))
    Q = R.orth(G, "QR")
    assert np.max(np.abs(Q.T @ Q - np.eye(4))) < 1e-13 and float(Q[:, 0] @ G[:, 0]) > 0
    P = R.orth(G, "SVD")
    assert np.max(np.abs(P.T @ P - np.eye(4))) < 1e-13
    S = P.T @ G
    assert np.max(np.abs(S - S.T)) < 1e-12 and np.all(np.linalg.eigvalsh((S + S.T) / 2) > 0)   # polar factor


@pytest.mark.parametrize("name", golden_cases())
def test_oracle_reproduces_golden(name):
    """The committed vectors are what oracle/ppls_ref.py computes today (guards silent edits)."""
    g = load_golden(name)
    i0 = init_of(g)
    out = R.ppls_simult_loop(g["X"], g["Y"], i0["W"], i0["C"], i0["B"], i0["sigE"], i0["sigF"], i0["sigH"], i0["sigT"],
                             EMsteps=int(g["EMsteps"]), atol=float(g["atol"]), type=str(g["type"]))
    assert len(out["loglik"]) == len(g["loglik"])
    assert relmax(out["loglik"], g["loglik"]) < 1e-12
    assert relmax(out["estimates"]["W"], g["est_W"]) < 1e-10
    assert relmax(np.diag(out["estimates"]["B"]), g["est_B"]) < 1e-10


def test_break_rule_golden_breaks_early():
    g = load_golden("c1_break_rule")
    n = len(g["loglik"])
    assert 2 < n < int(g["EMsteps"])
    assert g["loglik"][-1] - g["loglik"][-2] < float(g["atol"]) <= np.min(np.diff(g["loglik"][:-1]))


@pytest.mark.parametrize("name,nranks", [("c1_wellcond_1000", 1), ("c1_wellcond_1000", 3), ("c1_sig05_1000", 1),
                                         ("r5_odd_shapes", 2), ("r1_p_lt_q", 1), ("r8_max", 4), ("c1_break_rule", 1)])
def test_fused_algebra_matches_reference_restatement(oracle_lib, name, nranks):
    """The engine's algebra (one pass -> G1, G2, Gram; algebraic residuals; EMsteps+1 passes),
    run on the CPU with virtual row shards, reproduces the op-for-op restatement."""
    g = load_golden(name)
    i0 = init_of(g)
    f = F.fused_loop(oracle_lib, g["X"], g["Y"], i0["W"], i0["C"], i0["B"], i0["sigE"], i0["sigF"], i0["sigH"],
                     i0["sigT"], EMsteps=int(g["EMsteps"]), atol=float(g["atol"]), type=str(g["type"]), nranks=nranks)
    assert f["steps"] == len(g["loglik"])
    assert relmax(f["loglik"], g["loglik"]) < loglik_tol(name)
    for a_, b_ in ((f["W"], g["loop_W"]), (f["C"], g["loop_C"]), (f["B"], g["loop_B"]), (f["sigT"], g["loop_sigT"])):
        assert relmax(a_, b_) < TOL_PARAM
    for a_, b_ in ((f["Ctt"], g["Ctt"]), (f["Cuu"], g["Cuu"]), (f["Cut"], g["Cut"]), (f["Chh"], g["Chh"]),
                   (f["mu_T"], g["mu_T"]), (f["mu_U"], g["mu_U"])):
        assert relmax(a_, b_) < moment_tol(name)
    assert abs(f["Cee"] / g["Cee"] - 1) < TOL_PARAM and abs(f["Cff"] / g["Cff"] - 1) < TOL_PARAM


def test_reference_conditioning(oracle_lib):
    """Documents why the sigma = 0.05 case gets a looser log-likelihood tolerance: the op-for-op
    restatement itself moves by > 1e-11 relative under a 1-ulp perturbation of X."""
    g = load_golden("c1_sig05_1000")
    i0 = init_of(g)
    X2 = g["X"] * (1 + 1.2e-16 * np.sign(np.random.default_rng(0).standard_normal(g["X"].shape)))
    out = R.ppls_simult_loop(X2, g["Y"], i0["W"], i0["C"], i0["B"], i0["sigE"], i0["sigF"], i0["sigH"], i0["sigT"],
                             EMsteps=300, atol=-np.inf)
    d = np.max(np.abs(out["loglik"] - g["loglik"][:300]) / np.abs(g["loglik"][:300]))
    assert 1e-11 < d < 2e-9
    e = relmax(np.diag(out["Expectations"]["Ctt"]), np.diag(R.expect_m(
        g["X"], g["Y"], out["loop_state"]["W"], out["loop_state"]["C"], out["loop_state"]["B"], out["loop_state"]["sigE"],
        out["loop_state"]["sigF"], out["loop_state"]["sigH"], out["loop_state"]["sigT"])["Ctt"]))
    assert e < 1e-6     # same parameters, 1-ulp different X: the moment itself is this ill-conditioned


def test_generator_statistics_and_sharding(oracle_lib):
    """The counter-based generator gives the same bits for any row sharding, and N(0,1) moments."""
    truth = synth.truth_params(12, 6, 2)
    X, Y = cbind.gen_data(oracle_lib, 5, 0, 100, truth)
    Xb, Yb = cbind.gen_data(oracle_lib, 5, 40, 60, truth)
    assert np.array_equal(X[40:], Xb) and np.array_equal(Y[40:], Yb)
    z = cbind.gen_normals(oracle_lib, 9, 2, 0, 200000, 1)[:, 0]
    assert abs(z.mean()) < 0.01 and abs(z.std() - 1) < 0.01 and abs((z**4).mean() - 3) < 0.05
