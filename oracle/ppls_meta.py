"""CPU oracle for the multi-population ("meta") PPLS fit -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

numpy-FP64 restatement, operation for operation, of ("next" row 4)
  Package/PPLS/src/loglC.cpp:399-449   meta_Estep   -> meta_estep
  Package/PPLS/src/loglC.cpp:452-474   meta_Mstep   -> meta_mstep
  Package/PPLS/R/EM_W_multi.R:446-485  meta_EMstep  -> meta_emstep
  Package/PPLS/R/EM_W_multi.R:509-589  meta_PPLSi   -> meta_pplsi   (initialGuess "equal" / "custom", critfunc = identity,
                                                                    no constraints)
Rows must be sorted by population (the reference slices (1+Ni[i]):Ni[i+1], :455-456).  Quirks kept: each population's
Cxt, Cyu are divided by ITS OWN N before pooling (so populations weigh equally, :481-482), the sign of both pooled sums
comes from Cxt (:481-482), and logvalue[1,] holds the FULL-data log-likelihood once per population (:548), so the first
increment compares sum(logvalue[2,]) with nlevels times that value (:578).
PARITY: meta_estep / meta_mstep are PINNED to the reference's compiled meta_Estep / meta_Mstep (oracle/_ref,
tests/test_ref_pin.py); the R driver (meta_EMstep pooling, meta_PPLSi loop) is a restatement, unpinned.  Only tests/, smoke() and bench.py's CPU legs may import this module.
"""
from __future__ import annotations

import math
import numpy as np

from . import ppls_ref as R
from .ppls_seq import initial_guess


def meta_estep(W, C, B, X, Y, sigX, sigY, sigH, sigT, c1, c2, c3):
    sig2X, sig2Y, sig2H, sig2T = sigX * sigX, sigY * sigY, sigH * sigH, sigT * sigT
    N = X.shape[0]; p = W.size; q = C.size
    Xw = X @ W; Yc = Y @ C
    mu_T = Xw * sig2T * (-c1 + -c2 * B + 1 / sig2X) + Yc * sig2T * (-c2 + -c3 * B + 1 / sig2Y * B)
    Cxt = X.T @ mu_T / N
    Ctt = sig2T - sig2T * sig2T * (-c1 - 2 * B * c2 - B * B * (c3 - 1 / sig2Y) + 1 / sig2X) + float(mu_T @ mu_T) / N
    varU = sig2T * B * B + sig2H
    mu_U = (Xw * (-sig2T * B * c1 + -c2 * varU + 1 / sig2X * B * sig2T)
            + Yc * (-c2 * B * sig2T + -c3 * varU + 1 / sig2Y * varU))
    Cyu = Y.T @ mu_U / N
    Cut = sig2T * B - (-sig2T * sig2T * B * (c1 - 1 / sig2X) - sig2T * sig2T * B * B * c2 - sig2T * varU * c2
                       - varU * sig2T * B * (c3 - 1 / sig2Y)) + float(mu_U @ mu_T) / N
    xx, yy, xy = float(Xw @ Xw), float(Yc @ Yc), float(Xw @ Yc)
    Ceetmp = (c1 * c1 * sig2X * sig2X * xx + float(np.sum(X * X)) + c2 * c2 * sig2X * sig2X * yy
              - 2 * c1 * sig2X * xx + 2 * c1 * c2 * sig2X * sig2X * xy - 2 * c2 * sig2X * xy)
    Cee = sig2X - (-sig2X * sig2X * c1 + p * sig2X) / p + Ceetmp / N / p
    Cfftmp = (c3 * c3 * sig2Y * sig2Y * yy + float(np.sum(Y * Y)) + c2 * c2 * sig2Y * sig2Y * xx
              - 2 * c3 * sig2Y * yy + 2 * c3 * c2 * sig2Y * sig2Y * xy - 2 * c2 * sig2Y * xy)
    Cff = sig2Y - (-sig2Y * sig2Y * c3 + q * sig2Y) / q + Cfftmp / N / q
    mh = -c2 * sig2H * Xw - (c3 - 1 / sig2Y) * sig2H * Yc
    Chh = sig2H - (-sig2H * sig2H * (c3 - 1 / sig2Y)) + float(mh @ mh) / N
    return dict(Cxt=Cxt, Cyu=Cyu, mu_T=mu_T, mu_U=mu_U, Cut=Cut, Ctt=Ctt, Cee=Cee, Cff=Cff, Chh=Chh)


def meta_mstep(e):
    return dict(B=e["Cut"] / e["Ctt"], sighat=np.array([math.sqrt(e["Cee"]), math.sqrt(e["Cff"])]),
                siglathat=np.array([math.sqrt(e["Chh"]), math.sqrt(e["Ctt"])]), Cxt=e["Cxt"], Cyu=e["Cyu"])


def _unit(v):
    """orth() of one column: the normalised vector (QR of a single column with the reference's sign fix)."""
    return v / math.sqrt(float(v @ v))


def meta_emstep(X, Y, W, C, counts, params):
    Ni = np.concatenate([[0], np.cumsum(counts)])
    W = np.ravel(W); C = np.ravel(C)
    ret = []
    for i in range(len(counts)):
        Xi, Yi = X[Ni[i]:Ni[i + 1]], Y[Ni[i]:Ni[i + 1]]
        pr = params[i]
        B_T, sigX, sigY, sigH, sigT = pr["B_T"], pr["sigX"], pr["sigY"], pr["sigH"], pr["sigT"]
        g = sigT**2 * B_T**2 + sigH**2
        Kw = sigT**2 - sigT**4 * B_T**2 / sigY**2 + sigT**4 * B_T**2 * g / (sigY**2 * (g + sigY**2))
        Kc = g - sigT**4 * B_T**2 / sigX**2 + sigT**6 * B_T**2 / (sigX**2 * (sigT**2 + sigX**2))
        Kwc = (sigT**2 * B_T / (sigX**2 * sigY**2) - Kc * sigT**2 * B_T / (sigX**2 * sigY**2 * (Kc + sigY**2))
               - sigT**4 * B_T / (sigX**2 * sigY**2 * (sigT**2 + sigX**2))
               + Kc * sigT**4 * B_T / (sigX**2 * sigY**2 * (Kc + sigY**2) * (sigT**2 + sigX**2)))
        c1 = Kw / (sigX**2 * (Kw + sigX**2)); c3 = Kc / (sigY**2 * (Kc + sigY**2)); c2 = Kwc
        ret.append(meta_mstep(meta_estep(W, C, B_T, Xi, Yi, sigX, sigY, sigH, sigT, c1, c2, c3)))
    sgn = [float(np.sign(ret[0]["Cxt"] @ e["Cxt"])) for e in ret]
    Wn = _unit(sum(s * e["Cxt"] for s, e in zip(sgn, ret)))                      # :481
    Cn = _unit(sum(s * e["Cyu"] for s, e in zip(sgn, ret)))                      # :482
    return ret, Wn, Cn


def meta_pplsi(X, Y, counts, EMsteps=100, atol=1e-4, initialGuess="equal", customGuess=None):
    """counts: rows per population, in row order (table(Ipopu) for sorted Ipopu)."""
    X = np.asarray(X, dtype=np.float64); Y = np.asarray(Y, dtype=np.float64)
    N, p = X.shape; q = Y.shape[1]; J = len(counts)
    assert sum(counts) == N
    Wnw, Cnw, Bnw, signw, siglatnw = initial_guess(p, q, initialGuess, customGuess)
    logvalue = np.full((EMsteps + 1, J), np.nan)
    logvalue[0, :] = R.logl_w(X, Y, Wnw, Cnw, Bnw, signw[0], signw[1], siglatnw[0], siglatnw[1])        # :548
    params = [dict(B_T=Bnw, sigX=signw[0], sigY=signw[1], sigH=siglatnw[0], sigT=siglatnw[1]) for _ in range(J)]
    Ni = np.concatenate([[0], np.cumsum(counts)])
    i = 0
    for i in range(1, EMsteps + 1):
        fit, Wnw, Cnw = meta_emstep(X, Y, Wnw, Cnw, counts, params)
        params = [dict(B_T=f["B"], sigX=f["sighat"][0], sigY=f["sighat"][1], sigH=f["siglathat"][0], sigT=f["siglathat"][1])
                  for f in fit]
        for j in range(J):
            pr = params[j]
            logvalue[i, j] = R.logl_w(X[Ni[j]:Ni[j + 1]], Y[Ni[j]:Ni[j + 1]], Wnw, Cnw, pr["B_T"], pr["sigX"], pr["sigY"],
                                      pr["sigH"], pr["sigT"])                                          # :574-576
        if float(np.sum(logvalue[i, :])) - float(np.sum(logvalue[i - 1, :])) < atol:                     # :578
            break
    return dict(W=Wnw, C=Cnw, params=params, log=logvalue[1:i + 1, :], Number_steps=i)
