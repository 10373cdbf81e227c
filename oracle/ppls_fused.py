"""Fused single-pass CPU restatement of PPLS_simult -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Same algebra as the GPU engine (one pass over X and Y per EM iteration producing
G1 = X'mu_T, G2 = Y'mu_U, the 2r x 2r Gram of [Xw Yc] and nothing else; every second
moment, both residual variances and the log-likelihood then follow from those sums), but
computed on the host: the data pass by oracle/ppls_oracle.c::ppls_oracle_moments
(OpenMP), the r-sized update in numpy.  Used (a) to validate that algebra against the
op-for-op restatement ppls_ref.py without a GPU, including "virtual ranks" row sharding,
and (b) as the "best-effort fused CPU" baseline of bench.py.

Reference lines restated: Package/PPLS/R/EM_W_multi.R:672-680 (coefficients), :688-712
(E-step), :729-742 (M-step), :297-323 + Package/PPLS/src/loglC.cpp:331-336 (log-lik),
:780-793 (loop and break rule).  PARITY UNPINNED (see ppls_ref.py).
"""
from __future__ import annotations

import math
import numpy as np

from . import cbind
from . import ppls_ref as R


def posterior_coef(b, t, sigE, sigF, sigH):
    co = R.coefficients_estep(b, t, sigE, sigF, sigH)
    c1, c2, c3 = co["c1"], co["c2"], co["c3"]
    t2 = t**2
    varU = t2 * b**2 + sigH**2
    isE2, isF2 = sigE**-2, sigF**-2
    aT = isE2 * t2 - c1 * t2 - c2 * t2 * b
    bT = isF2 * (t2 * b) - c2 * t2 - c3 * b * t2
    aU = isE2 * (t2 * b) - c1 * t2 * b - c2 * varU
    bU = isF2 * varU - c2 * t2 * b - c3 * varU
    return co, aT, bT, aU, bU


def loglik_from_moments(mom, N, p, q, b, t, sigE, sigF, sigH):
    r = len(b)
    co = R.coefficients_logl(b, t, sigE, sigF, sigH)
    Gam = mom["Gam"]
    sE2, sF2 = sigE * sigE, sigF * sigF
    Logdiag = (float(np.sum(np.log(sE2 + t**2))) + (p - r) * math.log(sE2)
               + float(np.sum(np.log(sF2 + co["Kc"]))) + (q - r) * math.log(sF2))
    traceL = 1 / sE2 * mom["ssqX"] + 1 / sF2 * mom["ssqY"]
    for k in range(r):
        traceL += -co["c1"][k] * Gam[k, k] - 2 * co["c2"][k] * Gam[k, r + k] - co["c3"][k] * Gam[r + k, r + k]
    return -0.5 * N * (p + q) * math.log(2 * math.pi) - 0.5 * N * Logdiag - 0.5 * traceL


def estep_moments(mom, N, p, q, b, t, sigE, sigF, sigH, WtW, CtC):
    r = len(b)
    co, aT, bT, aU, bU = posterior_coef(b, t, sigE, sigF, sigH)
    c1, c2, c3 = co["c1"], co["c2"], co["c3"]
    Gam = mom["Gam"]
    Gxx, Gxy, Gyy = Gam[:r, :r], Gam[:r, r:], Gam[r:, r:]
    xx, xy, yy = np.diag(Gxx), np.diag(Gxy), np.diag(Gyy)
    t2, t4, b2 = t**2, t**4, b**2
    varU = t2 * b2 + sigH**2
    isE2, isF2 = sigE**-2, sigF**-2
    mTT = aT * aT * xx + 2 * aT * bT * xy + bT * bT * yy
    mUU = aU * aU * xx + 2 * aU * bU * xy + bU * bU * yy
    mUT = aU * aT * xx + (aU * bT + bU * aT) * xy + bU * bT * yy
    Ctt = np.abs(t2 - isE2 * t4 - isF2 * (t4 * b2) + t4 * c1 + 2 * t4 * b * c2 + t4 * b2 * c3 + mTT / N)
    Cuu = np.abs(varU - isE2 * (t4 * b2) - isF2 * varU**2 + t4 * b2 * c1 + 2 * t2 * b * varU * c2
                 + varU**2 * c3 + mUU / N)
    Cut = (t2 * b - isE2 * (t4 * b) - isF2 * (t2 * b * varU) + t4 * b * c1 + t2 * varU * c2 + t4 * b2 * c2
           + t2 * b * varU * c3 + mUT / N)
    sE2, sF2, sH2 = sigE**2, sigF**2, sigH**2
    A, Bm = sE2 * c1, sE2 * c2
    F, H = sF2 * c3, sF2 * c2
    MtM = np.outer(A, A) * Gxx + np.outer(A, Bm) * Gxy + np.outer(Bm, A) * Gxy.T + np.outer(Bm, Bm) * Gyy
    MtMf = np.outer(F, F) * Gyy + np.outer(F, H) * Gxy.T + np.outer(H, F) * Gxy + np.outer(H, H) * Gxx
    ssqE = mom["ssqX"] - 2 * float(np.sum(A * xx + Bm * xy)) + float(np.sum(MtM * WtW.T))
    ssqF = mom["ssqY"] - 2 * float(np.sum(F * yy + H * xy)) + float(np.sum(MtMf * CtC.T))
    Cee = (p * sE2 - p * sE2 + sigE**4 * float(np.sum(c1)) + ssqE / N) / p
    Cff = (q * sF2 - q * sF2 + sigF**4 * float(np.sum(c3)) + ssqF / N) / q
    hX, hY = -sH2 * c2, isF2 * sH2 - sH2 * c3
    Chh = (np.outer(hX, hX) * Gxx + np.outer(hX, hY) * Gxy + np.outer(hY, hX) * Gxy.T + np.outer(hY, hY) * Gyy) / N
    Chh = Chh + np.diag((sH2 - sigH**4 / sF2) + sigH**4 * c3)
    return dict(Ctt=Ctt, Cuu=Cuu, Cut=Cut, Cee=Cee, Cff=Cff, Chh=np.abs(Chh))


def fused_loop(L, X, Y, W0, C0, B0, sigE0, sigF0, sigH0, sigT0, EMsteps=10, atol=1e-4, type="SVD", nranks=1,
               want_mu=True, timings=None):
    """PPLS_simult from the init/loop seam on, with the engine's pass structure: pass k
    yields loglik(theta_k) and the E-step at theta_k; EMsteps + 1 passes when no break."""
    X = np.asfortranarray(X, dtype=np.float64); Y = np.asfortranarray(Y, dtype=np.float64)
    N, p = X.shape; q = Y.shape[1]
    W = np.array(W0, dtype=np.float64); Cm = np.array(C0, dtype=np.float64)
    b = R._diagvec(B0); t = R._diagvec(sigT0)
    sigE, sigF, sigH = float(sigE0), float(sigF0), float(sigH0)
    trace = []
    k = 0
    while True:
        _, aT, bT, aU, bU = posterior_coef(b, t, sigE, sigF, sigH)
        mom = cbind.moments(L, X, Y, W, Cm, aT, bT, aU, bU, nranks=nranks)
        ll = loglik_from_moments(mom, N, p, q, b, t, sigE, sigF, sigH)
        if k >= 1:
            trace.append(ll)
        E = estep_moments(mom, N, p, q, b, t, sigE, sigF, sigH, W.T @ W, Cm.T @ Cm)
        done = k >= EMsteps or (k >= 2 and trace[-1] - trace[-2] < atol)
        if done:
            break
        W = R.orth(mom["G1"], type); Cm = R.orth(mom["G2"], type)
        b = E["Cut"] * (1.0 / E["Ctt"]); t = np.sqrt(E["Ctt"])
        sigE = math.sqrt(E["Cee"]); sigF = math.sqrt(E["Cff"]); sigH = math.sqrt(float(np.sum(np.diag(E["Chh"]))) / len(b))
        k += 1
    out = dict(W=W, C=Cm, B=b, sigT=t, sigE=sigE, sigF=sigF, sigH=sigH, loglik=np.array(trace), steps=k, **E)
    if want_mu:
        _, aT, bT, aU, bU = posterior_coef(b, t, sigE, sigF, sigH)
        m2 = cbind.moments(L, X, Y, W, Cm, aT, bT, aU, bU, nranks=nranks, want_mu=True)
        out["mu_T"], out["mu_U"] = m2["mu_T"], m2["mu_U"]
    return out
