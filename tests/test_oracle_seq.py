"""CPU tests of the sequential-fit oracle (oracle/ppls_seq.py) against its committed golden vectors and
against properties the algorithm guarantees.  PARITY UNPINNED (no reference fixtures exist for this path)."""
import glob
import os

import numpy as np
import pytest

from helpers import GOLDEN_DIR, relmax
from oracle import ppls_ref, ppls_seq

SEQ_DIR = os.path.join(GOLDEN_DIR, "seq")


def seq_cases():
    return sorted(os.path.splitext(os.path.basename(f))[0] for f in glob.glob(os.path.join(SEQ_DIR, "*.npz")))


def load_seq(name):
    z = np.load(os.path.join(SEQ_DIR, name + ".npz"))
    g = {k: z[k] for k in z.files}
    g["guess"] = str(g["guess"])
    g["custom"] = None
    if g["guess"] == "custom":
        g["custom"] = {k[6:]: (g[k] if g[k].ndim else float(g[k])) for k in g if k.startswith("guess_")}
    return g


def test_there_are_golden_vectors():
    assert len(seq_cases()) >= 5


@pytest.mark.parametrize("name", seq_cases())
def test_oracle_reproduces_golden(name):
    g = load_seq(name)
    fit = ppls_seq.ppls(g["X"], g["Y"], int(g["a"]), int(g["EMsteps"]), float(g["atol"]), g["guess"], g["custom"])
    assert np.array_equal(fit["Number_steps"], g["Number_steps"])
    assert relmax(fit["W"], g["W"]) < 1e-12 and relmax(fit["C"], g["C"]) < 1e-12
    assert relmax(fit["B"], g["B"]) < 1e-12 and relmax(fit["sig"], g["sig"]) < 1e-12
    for i, t in enumerate(fit["logvalues"]):
        assert relmax(t, g["logvalues"][i, :t.size]) < 1e-13


@pytest.mark.parametrize("name", seq_cases())
def test_em_properties(name):
    """EM never decreases the likelihood; loadings of successive components are orthonormal (each is fitted
    on data deflated by the earlier ones, EM_W_multi.R:270-271)."""
    g = load_seq(name)
    k = g["W"].shape[1]
    for i in range(k):
        t = g["logvalues"][i, :int(g["Number_steps"][i]) + 1]
        assert np.all(np.diff(t) > -1e-8 * np.abs(t[:-1]))
    assert np.allclose(g["W"].T @ g["W"], np.eye(k), atol=1e-10)
    assert np.allclose(g["C"].T @ g["C"], np.eye(k), atol=1e-10)


def test_implicit_deflation_is_the_explicit_one():
    """The identity the GPU path rests on: for w orthogonal to the earlier loadings P, Xc w = X w and
    t(Xc) m = (I - P P') t(X) m, ssq(Xc) = ssq(X) - sum ||X P_j||^2 -- no deflated copy is needed."""
    g = load_seq("seq_equal_3")
    X, W = g["X"], g["W"]
    P = W[:, :2]
    X1 = X - (X @ P[:, :1]) @ P[:, :1].T
    Xc = X1 - (X1 @ P[:, 1:2]) @ P[:, 1:2].T
    w = W[:, 2]
    m = np.random.default_rng(0).standard_normal(X.shape[0])
    assert relmax(Xc @ w, X @ w) < 1e-12
    assert relmax(Xc.T @ m, (np.eye(X.shape[1]) - P @ P.T) @ (X.T @ m)) < 1e-12
    assert abs(np.sum(Xc * Xc) / (np.sum(X * X) - np.sum((X @ P[:, 0]) ** 2) - np.sum((X1 @ P[:, 1]) ** 2)) - 1) < 1e-12
    assert relmax(X1 @ P[:, 1], X @ P[:, 1]) < 1e-12


def test_first_step_is_emstep_of_the_simultaneous_oracle_at_r1():
    """EMstepC_fast (loglC.cpp:341-397) and Expect_M + Maximiz_M (EM_W_multi.R:637-742) are the same
    update for one component: two independent restatements of the reference agree."""
    g = load_seq("seq_p_lt_q_1")
    X, Y = g["X"], g["Y"]
    p, q = X.shape[1], Y.shape[1]
    W0, C0, B0, sg, sl = ppls_seq.initial_guess(p, q, "equal")
    one = ppls_seq.emstep_w(X, Y, W0, C0, B0, sg[0], sg[1], sl[0], sl[1])
    e = ppls_ref.expect_m(X, Y, W0[:, None], C0[:, None], np.array([[B0]]), sg[0], sg[1], sl[0], np.array([[sl[1]]]))
    m = ppls_ref.maximiz_m(e, X, Y, "SVD")
    assert relmax(m["W"][:, 0], one["W"]) < 1e-10 and relmax(m["C"][:, 0], one["C"]) < 1e-10
    assert abs(m["B"][0, 0] / one["B"] - 1) < 1e-10
    assert abs(m["sigE"] / one["sighat"][0] - 1) < 1e-10 and abs(m["sigF"] / one["sighat"][1] - 1) < 1e-10
    assert abs(m["sigH"] / one["siglathat"][0] - 1) < 1e-10 and abs(m["sigT"][0, 0] / one["siglathat"][1] - 1) < 1e-10
