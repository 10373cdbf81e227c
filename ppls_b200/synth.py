"""Generating parameters and start values for the synthetic PPLS workloads.

Host-side numpy on r-sized / (p+q) x r objects only; the data themselves are generated on
the device (ppls_data_synthetic).  The shapes of the true loadings follow the reference's
own simulation script Package/EM_Cpp.R:17-28 (orthonormalised Gaussian bumps,
B = 1.5 exp(-0.3 j), sigT = exp(-0.1 j), sigE = sigF = sigH = 0.05)."""
from __future__ import annotations

import numpy as np


def polar(G):
    u, _, vt = np.linalg.svd(np.asarray(G, dtype=np.float64), full_matrices=False)
    return u @ vt


def _bumps(n, r, centre0, width):
    x = np.arange(1, n + 1, dtype=np.float64)[:, None]
    j = np.arange(1, r + 1, dtype=np.float64)[None, :]
    mu = (centre0 + j / 10.0) * n
    sd = width * n
    M = np.exp(-0.5 * ((x - mu) / sd) ** 2) / (sd * np.sqrt(2 * np.pi))
    # keep the columns well separated for larger r than the reference script used
    return M + 1e-3 * np.cos(np.pi * (x - 0.5) * j / n) / np.sqrt(n)


def truth_params(p, q, r, sigE=0.05, sigF=0.05, sigH=0.05):
    W = polar(_bumps(p, r, 0.5 - 0.05 * r, 0.1))
    C = polar(_bumps(q, r, 0.6 - 0.05 * r, 0.1))
    j = np.arange(r, dtype=np.float64)
    return dict(W=W, C=C, B=1.5 * np.exp(-0.3 * j), sigT=np.exp(-0.1 * j), sigE=sigE, sigF=sigF, sigH=sigH)


def start_params(truth, seed=0, noise=0.5):
    """W0 = polar(W_true + noise Z/sqrt(p)), C0 likewise, B0 = I, sigT0 = I,
    sigE0 = 1/sqrt(p), sigF0 = 1/sqrt(q), sigH0 = 0.1 (replaces EM_W_multi.R:762-778)."""
    rng = np.random.default_rng(seed)
    W, C = np.asarray(truth["W"]), np.asarray(truth["C"])
    p, r = W.shape
    q = C.shape[0]
    return dict(W=polar(W + noise * rng.standard_normal((p, r)) / np.sqrt(p)),
                C=polar(C + noise * rng.standard_normal((q, r)) / np.sqrt(q)),
                B=np.ones(r), sigT=np.ones(r), sigE=1.0 / np.sqrt(p), sigF=1.0 / np.sqrt(q), sigH=0.1)
