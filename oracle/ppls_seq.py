"""CPU oracle for the sequential (one component at a time) PPLS fit -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

This is the path PPLS_simult uses to build its start values (Package/PPLS/R/EM_W_multi.R:762:
PPLS(X, Y, a, 20, 1e-4, 'random')) -- SURVEY.md section 8f, "next" row 1.  numpy-FP64 restatement,
operation for operation, of

  Package/PPLS/src/loglC.cpp:341-397   EMstepC_fast   -> emstepc_fast
  Package/PPLS/R/EM_W_multi.R:51-74    EMstep_W       -> emstep_w
  Package/PPLS/R/EM_W_multi.R:116-180  PPLSi          -> pplsi      (initialGuess "equal" / "custom";
                                                        critfunc = identity; no constraints)
  Package/PPLS/R/EM_W_multi.R:229-279  PPLS           -> ppls       (explicit deflation, :270-271)

"random" (:133, runif / rchisq from R's RNG) and "o2m" (OmicsPLS::o2m) start values are not restated:
R's RNG stream and OmicsPLS are not available here ("random" is restated distributionally in ppls_b200).
PARITY: emstepc_fast and the log-likelihood are PINNED to the reference's compiled EMstepC_fast / loglC_fast
(oracle/_ref, tests/test_ref_pin.py: single steps at 1e-12 and the whole golden fits replayed with the reference's
functions swapped in); the R driver around them (PPLSi / PPLS, deflation, break rule) is a restatement, unpinned.
Only tests/, smoke() and bench.py's CPU legs may import this module.
"""
from __future__ import annotations

import math
import numpy as np

from . import ppls_ref as R

EPS = float(np.finfo(np.float64).eps)


def emstepc_fast(W, C, B, X, Y, sigX, sigY, sigH, sigT, c1, c2, c3):
    """loglC.cpp:341-397 (r = 1): expectations and the maximisation that follows them."""
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
    Cuu = varU - (-(c1 - 1 / sig2X) * sig2T * sig2T * B * B - 2 * sig2T * B * varU * c2
                  - varU**2 * (c3 - 1 / sig2Y)) + float(mu_U @ mu_U) / N
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
    return dict(mu_T=mu_T, mu_U=mu_U, W=Cxt / math.sqrt(float(Cxt @ Cxt)), C=Cyu / math.sqrt(float(Cyu @ Cyu)),
                B=Cut / Ctt, sighat=np.array([math.sqrt(Cee), math.sqrt(Cff)]),
                siglathat=np.array([math.sqrt(Chh), math.sqrt(Ctt)]),
                Cut=Cut, Ctt=Ctt, Cuu=Cuu, Cee=Cee, Cff=Cff, Chh=Chh)


def emstep_w(X, Y, W, C, B, sigX, sigY, sigH, sigT):
    """EM_W_multi.R:51-74: closed-form coefficients (r = 1), then EMstepC_fast."""
    g = sigT**2 * B**2 + sigH**2
    Kw = sigT**2 - sigT**4 * B**2 / sigY**2 + sigT**4 * B**2 * g / (sigY**2 * (g + sigY**2))
    Kc = g - sigT**4 * B**2 / sigX**2 + sigT**6 * B**2 / (sigX**2 * (sigT**2 + sigX**2))
    Kwc = (sigT**2 * B / (sigX**2 * sigY**2) - Kc * sigT**2 * B / (sigX**2 * sigY**2 * (Kc + sigY**2))
           - sigT**4 * B / (sigX**2 * sigY**2 * (sigT**2 + sigX**2))
           + Kc * sigT**4 * B / (sigX**2 * sigY**2 * (Kc + sigY**2) * (sigT**2 + sigX**2)))
    c1 = Kw / (sigX**2 * (Kw + sigX**2)); c3 = Kc / (sigY**2 * (Kc + sigY**2)); c2 = Kwc
    return emstepc_fast(np.ravel(W), np.ravel(C), float(B), X, Y, sigX, sigY, sigH, sigT, c1, c2, c3)


def initial_guess(p, q, kind="equal", custom=None, seed=0, comp=0):
    """EM_W_multi.R:126-141 ("equal": orth(rep(1,p)) is the normalised ones vector; "random", :133: orth(runif(p)),
    orth(runif(q)), rchisq draws -- from the counter-based generator keyed by (seed, comp), the host restatement of
    ppls_b200's device kernel, since R's RNG stream does not exist outside R)."""
    if custom is not None or kind == "custom":
        g = custom
        return (np.ravel(g["W"]).astype(float), np.ravel(g["C"]).astype(float), float(g["B"]),
                np.array([float(g["sigE"]), float(g["sigF"])]), np.array([float(g["sigH"]), float(g["sigT"])]))
    if kind == "random":
        from . import cbind
        g = cbind.random_guess(cbind.load(), p, q, int(seed), int(comp))
        return (g["W"], g["C"], float(g["B"]), np.array([g["sigE"], g["sigF"]]), np.array([g["sigH"], g["sigT"]]))
    if kind != "equal":
        raise ValueError("oracle restates initialGuess 'equal', 'random' and 'custom' only")
    return (np.ones(p) / math.sqrt(p), np.ones(q) / math.sqrt(q), 1.0, np.array([1.0 / p, 1.0 / q]), np.array([1.0, 1.0]))


def _constrain(cons, Wnw, Cnw, Bnw, signw, siglatnw):
    """EM_W_multi.R:142-146 / :163-167: fixed values replace the estimates."""
    if not cons:
        return Wnw, Cnw, Bnw, signw, siglatnw
    g = lambda k: cons.get(k)                                                                    # noqa: E731
    Wnw = np.ravel(g("W")).astype(float) if g("W") is not None else Wnw
    Cnw = np.ravel(g("C")).astype(float) if g("C") is not None else Cnw
    Bnw = float(g("B")) if g("B") is not None else Bnw
    signw = np.array([float(g("sigE")) if g("sigE") is not None else signw[0], float(g("sigF")) if g("sigF") is not None else signw[1]])
    siglatnw = np.array([float(g("sigH")) if g("sigH") is not None else siglatnw[0],
                         float(g("sigT")) if g("sigT") is not None else siglatnw[1]])
    return Wnw, Cnw, Bnw, signw, siglatnw


def pplsi(X, Y, EMsteps=100, atol=1e-4, initialGuess="equal", customGuess=None, seed=0, comp=0, critfunc=None, constraints=None):
    """EM_W_multi.R:116-180; critfunc None = identity (:116), constraints a dict as fconstraint makes it (:85-92)."""
    crit = critfunc or (lambda x: x)
    N, p = X.shape; q = Y.shape[1]
    Wnw, Cnw, Bnw, signw, siglatnw = initial_guess(p, q, initialGuess, customGuess, seed, comp)
    Wnw, Cnw, Bnw, signw, siglatnw = _constrain(constraints, Wnw, Cnw, Bnw, signw, siglatnw)      # :142-146
    logvalue = [R.logl_w(X, Y, Wnw, Cnw, Bnw, signw[0], signw[1], siglatnw[0], siglatnw[1])]      # :150
    i = 0
    for i in range(1, EMsteps + 1):
        if np.any(signw < 100 * EPS):                                                               # :153-155
            return dict(W=None, C=None, B=None, sig=None, logvalue=None, Last_increment=None, Number_steps=i)
        fit = emstep_w(X, Y, Wnw, Cnw, Bnw, signw[0], signw[1], siglatnw[0], siglatnw[1])
        Bnw, Wnw, Cnw, signw, siglatnw = fit["B"], fit["W"], fit["C"], fit["sighat"], fit["siglathat"]
        Wnw, Cnw, Bnw, signw, siglatnw = _constrain(constraints, Wnw, Cnw, Bnw, signw, siglatnw)  # :163-167
        logvalue.append(R.logl_w(X, Y, Wnw, Cnw, Bnw, signw[0], signw[1], siglatnw[0], siglatnw[1]))  # :171
        if crit(logvalue[i] - logvalue[i - 1]) < atol:                                               # :172
            break
    return dict(W=Wnw, C=Cnw, B=Bnw, sig=np.concatenate([signw, siglatnw]), logvalue=np.array(logvalue),
                Last_increment=logvalue[i] - logvalue[i - 1], Number_steps=i)


def ppls(X, Y, nr_comp=1, EMsteps=100, atol=1e-4, initialGuess="equal", customGuess=None, seed=0, critfunc=None,
         constraints=None):
    """EM_W_multi.R:229-279: components one at a time with explicit deflation (:270-271)."""
    X = np.asarray(X, dtype=np.float64); Y = np.asarray(Y, dtype=np.float64)
    N, p = X.shape; q = Y.shape[1]; a = nr_comp
    Wn = np.full((p, a), np.nan); Cn = np.full((q, a), np.nan)
    Bn = np.full(a, np.nan); sig = np.full((a, 4), np.nan)
    last, nsteps, traces = [], [], []
    Xc, Yc = X.copy(), Y.copy()
    done = a
    for i in range(a):
        guess = customGuess[i] if isinstance(customGuess, (list, tuple)) else customGuess
        fit = pplsi(Xc, Yc, EMsteps, atol, initialGuess, guess, seed, i, critfunc, constraints[i] if constraints else None)
        if fit["B"] is None:                                                                         # :257-262
            done = i
            break
        Wn[:, i] = fit["W"]; Cn[:, i] = fit["C"]; Bn[i] = fit["B"]; sig[i] = fit["sig"]
        Xc = Xc - np.outer(Xc @ fit["W"], fit["W"])                                                   # :270
        Yc = Yc - np.outer(Yc @ fit["C"], fit["C"])                                                   # :271
        last.append(fit["Last_increment"]); nsteps.append(fit["Number_steps"]); traces.append(fit["logvalue"])
    return dict(W=Wn[:, :done], C=Cn[:, :done], B=Bn[:done], sig=sig[:done], Last_increment=np.array(last),
                Number_steps=np.array(nsteps, dtype=int), logvalues=traces)
