"""Multi-population ("meta") fit, meta_PPLSi (EM_W_multi.R:509-589; loglC.cpp:399-474): CPU checks of the oracle
restatement and GPU parity of ppls_meta_run / ppls_b200.meta_PPLSi against it.  PARITY UNPINNED."""
import numpy as np
import pytest

from helpers import relmax
from oracle import ppls_meta as M, ppls_seq


def pops_data(seed, counts, p, q):
    g = np.random.default_rng(seed)
    W = np.linalg.qr(g.standard_normal((p, 1)))[0]; C = np.linalg.qr(g.standard_normal((q, 1)))[0]
    Xs, Ys = [], []
    for j, n in enumerate(counts):
        T = g.standard_normal((n, 1)) * (1.5 + 0.3 * j)
        U = T * (1.2 - 0.15 * j) + 0.3 * g.standard_normal((n, 1))
        Xs.append(T @ W.T + (0.3 + 0.1 * j) * g.standard_normal((n, p)))
        Ys.append(U @ C.T + 0.4 * g.standard_normal((n, q)))
    return np.vstack(Xs), np.vstack(Ys), W[:, 0], C[:, 0]


def test_oracle_recovers_population_parameters():
    counts = [3000, 2000, 4500]
    X, Y, W, C = pops_data(1, counts, 18, 9)
    o = M.meta_pplsi(X, Y, counts, 80, 1e-6)
    assert abs(o["W"] @ W) > 0.999 and abs(o["C"] @ C) > 0.999
    for j, pr in enumerate(o["params"]):
        assert abs(abs(pr["B_T"]) - (1.2 - 0.15 * j)) < 0.05 and abs(pr["sigX"] - (0.3 + 0.1 * j)) < 0.02
    tot = o["log"].sum(axis=1)
    assert np.all(np.diff(tot) > -1e-8 * np.abs(tot[:-1]))


def test_one_population_first_step_is_the_sequential_step():
    """With one population meta_EMstep is EMstep_W: the two restatements agree (different reference functions)."""
    X, Y, _, _ = pops_data(2, [400], 12, 7)
    o = M.meta_pplsi(X, Y, [400], 1, -np.inf)
    s = ppls_seq.pplsi(X, Y, 1, -np.inf)
    assert relmax(o["W"], s["W"]) < 1e-12 and relmax(o["C"], s["C"]) < 1e-12
    assert abs(o["params"][0]["B_T"] / s["B"] - 1) < 1e-12 and abs(o["log"][0, 0] / s["logvalue"][1] - 1) < 1e-13


CASES = [([300, 200, 450], 18, 9, 40, 1e-4), ([257], 13, 7, 25, 1e-4), ([1000, 64, 333, 512, 90], 300, 77, 15, -np.inf),
         ([50, 60], 5, 30, 30, 1e-3)]


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
def test_gpu_meta_fit_matches_oracle(case):
    from ppls_b200 import meta_PPLSi
    counts, p, q, steps, atol = case
    X, Y, _, _ = pops_data(7 + p, counts, p, q)
    ref = M.meta_pplsi(X, Y, counts, steps, atol)
    lab = np.repeat(np.arange(len(counts)), counts)
    out = meta_PPLSi(X, Y, lab, steps, atol)
    assert out["Number_steps"] == ref["Number_steps"]
    assert relmax(out["W"], ref["W"]) < 1e-8 and relmax(out["C"], ref["C"]) < 1e-8
    for a, b in zip(out["params"], ref["params"]):
        for key in ("B_T", "sigX", "sigY", "sigH", "sigT"):
            assert abs(a[key] / b[key] - 1) < 1e-8, key
    assert out["log"].shape == ref["log"].shape and relmax(out["log"], ref["log"]) < 1e-10


@pytest.mark.gpu
def test_gpu_meta_fit_custom_guess_and_errors():
    from ppls_b200 import meta_PPLSi
    counts = [120, 80]
    X, Y, _, _ = pops_data(3, counts, 10, 6)
    g = np.random.default_rng(0)
    w = g.standard_normal(10); c = g.standard_normal(6)
    guess = dict(W=w / np.linalg.norm(w), C=c / np.linalg.norm(c), B=0.8, sigE=0.7, sigF=0.9, sigH=0.5, sigT=1.2)
    ref = M.meta_pplsi(X, Y, counts, 12, 1e-4, "custom", guess)
    out = meta_PPLSi(X, Y, np.repeat([0, 1], counts), 12, 1e-4, customGuess=guess)
    assert relmax(out["W"], ref["W"]) < 1e-8 and relmax(out["log"], ref["log"]) < 1e-10
    with pytest.raises(ValueError):
        meta_PPLSi(X, Y, np.array([0, 1] * 100), 5)          # populations not contiguous
    with pytest.raises(ValueError):
        meta_PPLSi(X, Y, np.repeat([0, 1], [100, 99]), 5)    # length(Ipopu) != nrow(X)
