"""Golden vectors for the sequential PPLS fit (oracle/ppls_seq.py::ppls), written to tests/golden/seq/.

PARITY UNPINNED: the reference holds no fixtures for this path and R is not installed here; the vectors
are outputs of the op-for-op numpy restatement.  Run from the repo root: python tests/golden/make_golden_seq.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ppls_seq  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "seq")


def model_data(seed, n, p, q, r, noise):
    """PO2PLS-style data: orthonormal loadings, decaying latent scales, isotropic noise."""
    g = np.random.default_rng(seed)
    W = np.linalg.qr(g.standard_normal((p, r)))[0]
    Cm = np.linalg.qr(g.standard_normal((q, r)))[0]
    sT = 2.0 * np.exp(-0.3 * np.arange(r))
    B = 1.5 - 0.2 * np.arange(r)
    T = g.standard_normal((n, r)) * sT
    U = T * B + 0.3 * g.standard_normal((n, r))
    X = T @ W.T + noise * g.standard_normal((n, p))
    Y = U @ Cm.T + noise * g.standard_normal((n, q))
    return X, Y


CASES = {
    # name: (seed, n, p, q, a, EMsteps, atol, noise, guess)
    "seq_equal_3": (11, 400, 24, 10, 3, 60, 1e-4, 0.3, "equal"),
    "seq_odd_2": (12, 257, 13, 7, 2, 40, 1e-6, 0.5, "equal"),
    "seq_allsteps_2": (13, 300, 16, 16, 2, 25, -np.inf, 0.4, "equal"),
    "seq_custom_2": (14, 350, 40, 9, 2, 50, 1e-5, 0.3, "custom"),
    "seq_break_2": (16, 500, 30, 12, 2, 100, 5e-2, 0.3, "equal"),
    "seq_p_lt_q_1": (15, 200, 5, 30, 1, 30, 1e-4, 0.2, "equal"),
}


def custom_guess(seed, p, q):
    g = np.random.default_rng(seed + 1000)
    w = g.standard_normal(p); c = g.standard_normal(q)
    return dict(W=w / np.linalg.norm(w), C=c / np.linalg.norm(c), B=0.7, sigE=0.9, sigF=1.1, sigH=0.4, sigT=1.3)


def main():
    os.makedirs(OUT, exist_ok=True)
    for name, (seed, n, p, q, a, steps, atol, noise, guess) in CASES.items():
        X, Y = model_data(seed, n, p, q, max(a, 2), noise)
        cg = custom_guess(seed, p, q) if guess == "custom" else None
        fit = ppls_seq.ppls(X, Y, a, steps, atol, guess, cg)
        k = fit["W"].shape[1]
        logv = np.full((k, steps + 1), np.nan)
        for i, t in enumerate(fit["logvalues"]):
            logv[i, :t.size] = t
        extra = {}
        if cg is not None:
            extra = {"guess_" + key: np.asarray(val, dtype=np.float64) for key, val in cg.items()}
        np.savez_compressed(os.path.join(OUT, name + ".npz"), X=X, Y=Y, a=a, EMsteps=steps, atol=atol, guess=guess,
                            W=fit["W"], C=fit["C"], B=fit["B"], sig=fit["sig"], Last_increment=fit["Last_increment"],
                            Number_steps=fit["Number_steps"], logvalues=logv, **extra)
        print(name, "steps", fit["Number_steps"], "B", fit["B"])


if __name__ == "__main__":
    main()
