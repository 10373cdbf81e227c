"""CPU oracle for variances.PPLS_simult -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

numpy-FP64 restatement, operation for operation, of Package/PPLS/R/EM_W_multi.R:830-860 (SURVEY.md section 8f,
"next" row 2): asymptotic standard errors of the loadings from the Expectations a PPLS_simult fit returns.
`diag(c(Ctt), nrow(X))` (:846) is an N x N diagonal matrix in the reference; scaling the rows of X by the
scalar is the same product.  PARITY UNPINNED (no fixtures in the reference, R absent; see oracle/ppls_ref.py).
Only tests/, smoke() and bench.py's CPU legs may import this module.
"""
from __future__ import annotations

import numpy as np

from . import ppls_ref as R


def variances_ppls_simult(fit, data, XorY="X"):
    """fit: the list PPLS_simult returns ({'Expectations': {mu_T, mu_U, Ctt, Cuu, ...}, 'estimates': {sigE, ...}}).
    Returns {'components': [{B_exp, SSt_exp, SSt_star}, ...], 'varMatrix': [...], 'seLoad': w x a}."""
    X = np.asarray(data, dtype=np.float64)
    Ex, est = fit["Expectations"], fit["estimates"]
    if XorY == "X":
        W = R.orth(X.T @ Ex["mu_T"], type="SVD")                              # :831
    elif XorY == "Y":
        W = R.orth(X.T @ Ex["mu_U"], type="SVD")                              # :832
    else:
        raise ValueError("XorY must be 'X' or 'Y'")
    N = X.shape[0]
    W = W.reshape(X.shape[1], -1)
    a = W.shape[1]
    Cmat = np.asarray(Ex["Ctt" if XorY == "X" else "Cuu"], dtype=np.float64)  # :836
    mu = np.asarray(Ex["mu_T" if XorY == "X" else "mu_U"], dtype=np.float64)  # :837
    sigE = float(est["sigE"])
    comps = []
    for i in range(a):
        w = W[:, [i]]                                                         # :839
        Ctt = N * Cmat[i, i]                                                  # :840
        m = mu[:, i]                                                          # :841
        cp = float(m @ m)
        Vt = Ctt - cp                                                         # :842
        Cxt = X.T @ m[:, None]                                                # :843
        B_star = Ctt / sigE**2 * np.eye(X.shape[1]) / N                       # :844
        SSt = (X.T @ (Ctt * X) - Cxt * (Ctt + 2 * Vt) @ w.T - w * (Ctt + 2 * Vt) @ Cxt.T
               + w * (Ctt**2 + 4 * cp * Vt + 2 * Vt * Vt) @ w.T)              # :846-847
        SSt = SSt / sigE**4 / N                                               # :848
        d = Cxt - w * Ctt
        SStar = (d @ d.T) / sigE**4                                           # :850-851
        comps.append(dict(B_exp=B_star, SSt_exp=SSt, SSt_star=SStar))
    varM = [-np.linalg.solve(c["B_exp"] - c["SSt_exp"], np.eye(X.shape[1])) for c in comps]   # :855
    se = np.stack([np.sqrt(np.diag(v)) for v in varM], axis=1)                # :856
    return dict(components=comps, varMatrix=varM, seLoad=se, W=W)
