"""CPU oracle for the PPLS_simult hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

numpy-FP64 restatement, operation for operation, of the reference's simultaneous
EM estimator (selbouhaddani/PPLS).  Only tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py may import this module; the
product path (ppls_b200/) never does.

PARITY: PARTLY PINNED.  The reference ships no tests, golden vectors or recorded output for
PPLS_simult and R is not installed here, but its one native file,
Package/PPLS/src/loglC.cpp, IS compiled unmodified into oracle/_ref/libppls_ref.so (stand-in
Eigen/Rcpp headers, oracle/ref_shim/; recipe `make -C oracle ref`), and tests/test_ref_pin.py holds
this module to it: loglc_fast == loglC_fast (1e-13, every golden case, whole traces replayed),
logl_w == the dense Cholesky log-likelihood of loglC (1e-9), expect_m's moments == the dense
E-step EMstepC (1e-9, one component).  UNPINNED (R code, no interpreter; `orth`/`ssq` live in the
un-vendored, un-versioned OmicsPLS package): the R-level expression order of Expect_M / Maximiz_M,
the normalisation quirk and orth -- defended by independent cross-checks instead (dense `debug=TRUE`
E-step, dense Gaussian log-likelihood, closed-form inverse identity: tests/test_oracle.py).
Golden vectors: tests/golden/, made by tests/golden/make_golden.py from this module.

Reference files restated (paths relative to /root/reference/Package/PPLS):
  R/EM_W_multi.R:758-807   PPLS_simult      -> ppls_simult_loop / normalise
  R/EM_W_multi.R:637-717   Expect_M         -> expect_m (fast) / expect_m_dense (debug)
  R/EM_W_multi.R:729-742   Maximiz_M        -> maximiz_m
  R/EM_W_multi.R:297-323   logl_W           -> logl_w
  src/loglC.cpp:319-338    loglC_fast       -> loglc_fast
  R/EM_W_multi.R:606-618   sseXY_W          -> sse_xy_w
  R/PJSC.R:1-5             tr               -> tr
  OmicsPLS::orth, ::ssq    (historical copy Package/functions.R:252-260, :380-385)
"""
from __future__ import annotations

import math
import numpy as np

__all__ = [
    "tr", "ssq", "orth", "coefficients_estep", "coefficients_logl", "expect_m",
    "expect_m_dense", "maximiz_m", "logl_w", "loglc_fast", "sse_xy_w",
    "normalise", "ppls_simult_loop", "dense_loglik",
]


def tr(X):
    """R/PJSC.R:1-5 -- trace (the non-square warning is irrelevant here)."""
    X = np.atleast_2d(np.asarray(X, dtype=np.float64))
    return float(np.sum(np.diag(X)))


def ssq(X):
    """OmicsPLS::ssq == sum(X^2) (Package/functions.R:380-385)."""
    X = np.asarray(X, dtype=np.float64)
    return float(np.sum(X * X))


def orth(G, type="SVD"):
    """OmicsPLS::orth (historical in-tree copy Package/functions.R:252-260).

    type "SVD": svd(X) -> tcrossprod(u, v)  (the polar factor, LAPACK dgesdd in R
    and in numpy).  type "QR": qr.Q(qr(X)) times one global sign that makes column 1
    positively correlated with X[,1].  LINPACK dqrdc2 (R) and LAPACK dgeqrf (numpy)
    share the Householder sign convention R_ll = -sign(x_ll)*||x_l||, so Q agrees
    whenever dqrdc2 does not pivot (full-column-rank input).
    """
    G = np.asarray(G, dtype=np.float64)
    if G.ndim == 1:
        G = G[:, None]
    if type == "SVD":
        u, _, vt = np.linalg.svd(G, full_matrices=False)
        return u @ vt
    if type == "QR":
        e, _ = np.linalg.qr(G)
        s = np.sign(float(e[:, 0] @ G[:, 0]))
        return s * e
    raise ValueError("type must be 'SVD' or 'QR'")


def coefficients_estep(b, t, sigE, sigF, sigH):
    """Closed-form inverse coefficients as Expect_M evaluates them
    (R/EM_W_multi.R:672-680).  b, t: length-r diagonals of B and sigT."""
    b = np.asarray(b, dtype=np.float64)
    t = np.asarray(t, dtype=np.float64)
    a = len(b)
    sigE = float(sigE); sigF = float(sigF); sigH = float(sigH)
    g = np.empty(a); Kw = np.empty(a); Kc = np.empty(a); Kwc = np.empty(a)
    c1 = np.empty(a); c3 = np.empty(a)
    for i in range(a):
        ti = float(t[i]); bi = float(b[i])
        g[i] = ti**2 * bi**2 + sigH**2
        gi = float(g[i])
        Kw[i] = ti**2 - ti**4 * bi**2 / sigF**2 + ti**4 * bi**2 * gi / (sigF**2 * (gi + sigF**2))
        Kc[i] = gi - ti**4 * bi**2 / sigE**2 + ti**6 * bi**2 / (sigE**2 * (ti**2 + sigE**2))
        kc = float(Kc[i])
        Kwc[i] = (ti**2 * bi / (sigE**2 * sigF**2)
                  - kc * ti**2 * bi / (sigE**2 * sigF**2 * (kc + sigF**2))
                  - ti**4 * bi / (sigE**2 * sigF**2 * (ti**2 + sigE**2))
                  + kc * ti**4 * bi / (sigE**2 * sigF**2 * (kc + sigF**2) * (ti**2 + sigE**2)))
        c1[i] = Kw[i] / (sigE**2 * (Kw[i] + sigE**2))
        c3[i] = Kc[i] / (sigF**2 * (Kc[i] + sigF**2))
    c2 = Kwc.copy()
    return dict(g=g, Kw=Kw, Kc=Kc, Kwc=Kwc, c1=c1, c2=c2, c3=c3)


def coefficients_logl(b, t, sigX, sigY, sigH):
    """Same coefficients as logl_W evaluates them (R/EM_W_multi.R:312-320):
    g is stored as a square ROOT there and squared again on use."""
    b = np.asarray(b, dtype=np.float64)
    t = np.asarray(t, dtype=np.float64)
    a = len(b)
    sigX = float(sigX); sigY = float(sigY); sigH = float(sigH)
    g = np.empty(a); Kw = np.empty(a); Kc = np.empty(a); Kwc = np.empty(a)
    c1 = np.empty(a); c3 = np.empty(a)
    for i in range(a):
        ti = float(t[i]); bi = float(b[i])
        g[i] = math.sqrt(ti**2 * bi**2 + sigH**2)
        gi = float(g[i])
        Kw[i] = ti**2 - ti**4 * bi**2 / sigY**2 + ti**4 * bi**2 * gi**2 / (sigY**2 * (gi**2 + sigY**2))
        Kc[i] = gi**2 - ti**4 * bi**2 / sigX**2 + ti**6 * bi**2 / (sigX**2 * (ti**2 + sigX**2))
        kc = float(Kc[i])
        Kwc[i] = (ti**2 * bi / (sigX**2 * sigY**2)
                  - kc * ti**2 * bi / (sigX**2 * sigY**2 * (kc + sigY**2))
                  - ti**4 * bi / (sigX**2 * sigY**2 * (ti**2 + sigX**2))
                  + kc * ti**4 * bi / (sigX**2 * sigY**2 * (kc + sigY**2) * (ti**2 + sigX**2)))
        c1[i] = Kw[i] / (sigX**2 * (Kw[i] + sigX**2))
        c3[i] = Kc[i] / (sigY**2 * (Kc[i] + sigY**2))
    c2 = Kwc.copy()
    return dict(g=g, Kw=Kw, Kc=Kc, Kwc=Kwc, c1=c1, c2=c2, c3=c3)


def _diagvec(M, a=None):
    """Accept an r-vector, a scalar or an r x r (diagonal) matrix; return the r-vector."""
    M = np.asarray(M, dtype=np.float64)
    if M.ndim == 2:
        return np.diag(M).copy()
    return np.atleast_1d(M).astype(np.float64).copy()


def sse_xy_w(W, C, B, sigX, sigY, sigH, sigT):
    """R/EM_W_multi.R:606-618 -- the dense (p+q) x (p+q) model covariance."""
    W = np.asarray(W, dtype=np.float64); C = np.asarray(C, dtype=np.float64)
    p = W.shape[0]; q = C.shape[0]
    Bm = np.diag(_diagvec(B)); Tm = np.diag(_diagvec(sigT))
    WT = W @ Tm
    SX = WT @ WT.T + sigX**2 * np.eye(p)
    SXY = W @ Bm @ (Tm**2) @ C.T
    CBT = C @ Bm.T @ Tm
    SY = CBT @ CBT.T + (C @ C.T) * sigH**2 + sigY**2 * np.eye(q)
    return np.block([[SX, SXY], [SXY.T, SY]])


def expect_m(X, Y, W, C, B, sigE, sigF, sigH, sigT):
    """Expect_M, fast branch (R/EM_W_multi.R:668-713, return :715-716).

    Written with the same temporaries and the same left-to-right products of
    diagonal r x r matrices as the R source, so that this function is also the
    'reference-faithful' timed CPU baseline (numpy -> BLAS as R %*% -> BLAS)."""
    X = np.asarray(X, dtype=np.float64); Y = np.asarray(Y, dtype=np.float64)
    W = np.asarray(W, dtype=np.float64); C = np.asarray(C, dtype=np.float64)
    N = X.shape[0]; p = X.shape[1]; q = Y.shape[1]; a = W.shape[1]
    sigE = float(sigE); sigF = float(sigF); sigH = float(sigH)
    tvec = _diagvec(sigT); bvec = _diagvec(B)
    co = coefficients_estep(bvec, tvec, sigE, sigF, sigH)
    sigT = np.diag(tvec); B = np.diag(bvec)
    c1 = np.diag(co["c1"]); c2 = np.diag(co["c2"]); c3 = np.diag(co["c3"])
    I = np.eye(a)
    sigT2 = sigT**2; sigT4 = sigT**4; B2 = B**2

    varU = sigT2 @ B2 + sigH**2 * I                                         # :688
    Xw = X @ W                                                              # :689
    Yc = Y @ C                                                              # :690
    mu_T = (sigE**-2 * (Xw @ sigT2) + sigF**-2 * (Yc @ sigT2 @ B) - Xw @ c1 @ sigT2
            - Xw @ c2 @ sigT2 @ B - Yc @ c2 @ sigT2 - Yc @ c3 @ B @ sigT2)   # :691-692
    mu_U = (sigE**-2 * (Xw @ sigT2 @ B) + sigF**-2 * (Yc @ varU)
            - Xw @ c1 @ sigT2 @ B - Xw @ c2 @ varU - Yc @ c2 @ sigT2 @ B - Yc @ c3 @ varU)  # :693-694

    Ctt = (sigT2 - sigE**-2 * sigT4 - sigF**-2 * (sigT4 @ B2) + sigT4 @ c1 + 2 * sigT4 @ B @ c2
           + sigT4 @ B2 @ c3 + (mu_T.T @ mu_T) / N)                         # :696-697
    Cuu = (varU - sigE**-2 * (sigT4 @ B2) - sigF**-2 * varU**2 + sigT4 @ B2 @ c1
           + 2 * sigT2 @ B @ varU @ c2 + varU**2 @ c3 + (mu_U.T @ mu_U) / N)  # :698-699
    Cut = (sigT2 @ B - sigE**-2 * (sigT4 @ B) - sigF**-2 * (sigT2 @ B @ varU) + sigT4 @ B @ c1
           + sigT2 @ varU @ c2 + sigT4 @ B2 @ c2 + sigT2 @ B @ varU @ c3
           + (mu_U.T @ mu_T) / N)                                           # :700-701

    mu_E = X - sigE**2 * (Xw @ c1 @ W.T) - sigE**2 * (Yc @ c2 @ W.T)        # :703
    Cee = (p * sigE**2 - p * sigE**2 + sigE**4 * np.sum(c1) + ssq(mu_E) / N) / p   # :706
    mu_F = Y - sigF**2 * (Yc @ c3 @ C.T) - sigF**2 * (Xw @ c2 @ C.T)        # :708
    Cff = (q * sigF**2 - q * sigF**2 + sigF**4 * np.sum(c3) + ssq(mu_F) / N) / q   # :709
    mu_H = sigF**-2 * sigH**2 * Yc - sigH**2 * (Xw @ c2 + Yc @ c3)          # :711
    Chh = (sigH**2 - sigH**4 / sigF**2) * I + sigH**4 * c3 + (mu_H.T @ mu_H) / N   # :712

    return dict(mu_T=mu_T, mu_U=mu_U, Ctt=np.abs(Ctt) * I, Cuu=np.abs(Cuu) * I, Cut=Cut * I,
                Cee=np.array([[Cee]]), Cff=np.array([[Cff]]), Chh=np.abs(Chh))     # :715-716


def expect_m_dense(X, Y, W, C, B, sigE, sigF, sigH, sigT):
    """Expect_M(debug=TRUE) (R/EM_W_multi.R:643-667): dense solve() form, the
    reference's own slow cross-check.  O((p+q)^3); small shapes only."""
    X = np.asarray(X, dtype=np.float64); Y = np.asarray(Y, dtype=np.float64)
    W = np.asarray(W, dtype=np.float64); C = np.asarray(C, dtype=np.float64)
    N = X.shape[0]; p = X.shape[1]; q = Y.shape[1]; a = W.shape[1]
    sigT = np.diag(_diagvec(sigT)); B = np.diag(_diagvec(B)); I = np.eye(a)
    covT = np.vstack([W @ sigT**2, C @ B @ sigT**2])
    covU = np.vstack([W @ B @ sigT**2, C @ B**2 @ sigT**2 + sigH**2 * C])
    invS = np.linalg.inv(sse_xy_w(W, C, B, sigE, sigF, sigH, sigT))
    XY = np.hstack([X, Y])
    mu_T = XY @ invS @ covT
    mu_U = XY @ invS @ covU
    sigU2 = sigT**2 @ B**2 + sigH**2 * I
    Ctt = sigT**2 - covT.T @ invS @ covT + (mu_T.T @ mu_T) / N
    Cuu = sigU2 - covU.T @ invS @ covU + (mu_U.T @ mu_U) / N
    Cut = sigT**2 @ B - covU.T @ invS @ covT + (mu_U.T @ mu_T) / N
    covE = np.vstack([sigE**2 * np.eye(p), np.zeros((q, p))])
    mu_E = XY @ invS @ covE
    Cee = tr(sigE**2 * np.eye(p) - covE.T @ invS @ covE + (mu_E.T @ mu_E) / N) / p
    covF = np.vstack([np.zeros((p, q)), sigF**2 * np.eye(q)])
    mu_F = XY @ invS @ covF
    Cff = tr(sigF**2 * np.eye(q) - covF.T @ invS @ covF + (mu_F.T @ mu_F) / N) / q
    covH = np.vstack([0 * W, sigH**2 * C])
    mu_H = XY @ invS @ covH
    Chh = sigH**2 * I - covH.T @ invS @ covH + (mu_H.T @ mu_H) / N
    return dict(mu_T=mu_T, mu_U=mu_U, Ctt=np.abs(Ctt) * I, Cuu=np.abs(Cuu) * I, Cut=Cut * I,
                Cee=np.array([[Cee]]), Cff=np.array([[Cff]]), Chh=np.abs(Chh),
                Ctt_full=Ctt, Cuu_full=Cuu, Cut_full=Cut)


def maximiz_m(fit, X, Y, type="SVD"):
    """Maximiz_M (R/EM_W_multi.R:729-742)."""
    X = np.asarray(X, dtype=np.float64); Y = np.asarray(Y, dtype=np.float64)
    Cut = fit["Cut"]; Ctt = fit["Ctt"]; a = Cut.shape[0]
    I = np.eye(a)
    return dict(
        W=orth(X.T @ fit["mu_T"], type=type),                   # :732 (explicit t(X) in R)
        C=orth(Y.T @ fit["mu_U"], type=type),                   # :733
        B=(Cut @ np.linalg.inv(Ctt)) * I,                       # :734  Cut %*% solve(Ctt) * I
        sigE=math.sqrt(tr(fit["Cee"]) / fit["Cee"].shape[1]),   # :735
        sigF=math.sqrt(tr(fit["Cff"]) / fit["Cff"].shape[1]),   # :736
        sigH=math.sqrt(tr(fit["Chh"]) / fit["Chh"].shape[1]),   # :737
        sigT=np.sqrt(Ctt * I),                                  # :738
    )


def loglc_fast(W, C, X, Y, sigX, sigY, sig2T, c1, c2, c3, Kc):
    """loglC_fast (src/loglC.cpp:319-338; the src/ copy is authoritative)."""
    W = np.asarray(W, dtype=np.float64); C = np.asarray(C, dtype=np.float64)
    X = np.asarray(X, dtype=np.float64); Y = np.asarray(Y, dtype=np.float64)
    sig2X = sigX * sigX
    sig2Y = sigY * sigY
    N = X.shape[0]; p = W.shape[0]; q = C.shape[0]; a = W.shape[1]
    Logdiag = (float(np.sum(np.log(sig2X + sig2T))) + (p - a) * math.log(sig2X)
               + float(np.sum(np.log(sig2Y + Kc))) + (q - a) * math.log(sig2Y))          # :331
    XW = X @ W                                                                           # :332
    YC = Y @ C                                                                           # :333
    traceL = 1 / sig2X * float(np.sum(X * X)) + 1 / sig2Y * float(np.sum(Y * Y))         # :334
    for i in range(a):                                                                   # :335
        traceL += (-c1[i] * float(XW[:, i] @ XW[:, i]) - 2 * c2[i] * float(XW[:, i] @ YC[:, i])
                   - c3[i] * float(YC[:, i] @ YC[:, i]))
    return -0.5 * N * (p + q) * math.log(2 * math.pi) - 0.5 * N * Logdiag - 0.5 * traceL  # :336


def logl_w(X, Y, W, C, B, sigX, sigY, sigH, sigT):
    """logl_W (R/EM_W_multi.R:297-323)."""
    W = np.asarray(W, dtype=np.float64); C = np.asarray(C, dtype=np.float64)
    if W.ndim == 1: W = W[:, None]
    if C.ndim == 1: C = C[:, None]
    tvec = _diagvec(sigT); bvec = _diagvec(B)
    co = coefficients_logl(bvec, tvec, sigX, sigY, sigH)
    return loglc_fast(W, C, X, Y, float(sigX), float(sigY), tvec**2,
                      co["c1"], co["c2"], co["c3"], co["Kc"])


def dense_loglik(X, Y, W, C, B, sigX, sigY, sigH, sigT):
    """Independent cross-check: Gaussian log-likelihood of the rows of [X Y] under the
    dense covariance sseXY_W (cf. Package/functions.R:198-227 logl_M, plus the 2*pi
    constant the src/ copy of loglC_fast includes)."""
    S = sse_xy_w(W, C, B, sigX, sigY, sigH, sigT)
    XY = np.hstack([np.asarray(X, dtype=np.float64), np.asarray(Y, dtype=np.float64)])
    N, d = XY.shape
    _, logdet = np.linalg.slogdet(S)
    quad = float(np.sum((XY @ np.linalg.inv(S)) * XY))
    return -0.5 * N * d * math.log(2 * math.pi) - 0.5 * N * logdet - 0.5 * quad


def normalise(W, C, B, sigT):
    """Sign/order normalisation used on the start values (R/EM_W_multi.R:773-778)
    and on the result (:794-799).  Reproduces the quirk that for W and C the sign
    vector is applied AFTER the column permutation without being permuted itself,
    whereas for B it is applied before the permutation."""
    W = np.asarray(W, dtype=np.float64); C = np.asarray(C, dtype=np.float64)
    b = _diagvec(B); t = _diagvec(sigT)
    signLoad = np.sign(t * b)                                   # sign(diag(sigT %*% B))
    key = t * b * signLoad
    rotLoad = np.argsort(-key, kind="stable")                   # order(..., decreasing=TRUE)
    Wn = W[:, rotLoad] @ np.diag(signLoad)
    Cn = C[:, rotLoad] @ np.diag(signLoad)
    Bn = np.diag((b * signLoad)[rotLoad])
    Tn = np.diag(t[rotLoad])
    return Wn, Cn, Bn, Tn


def ppls_simult_loop(X, Y, W0, C0, B0, sigE0, sigF0, sigH0, sigT0, EMsteps=10, atol=1e-4,
                     type="SVD", debug=False, record=None):
    """PPLS_simult from the init/loop seam on (R/EM_W_multi.R:780-806).

    The reference draws its start values at random (:762-778); parity
    needs 'same inputs and initialisation', so the start state is an argument here.
    Returns the same list structure: Expectations, loglik, estimates.
    `record(i, state)` is called after every EM step with the loop-order state."""
    X = np.asarray(X, dtype=np.float64); Y = np.asarray(Y, dtype=np.float64)
    a = np.asarray(W0).shape[1]
    W_ = np.array(W0, dtype=np.float64); C_ = np.array(C0, dtype=np.float64)
    B_ = np.diag(_diagvec(B0)); sigT_ = np.diag(_diagvec(sigT0))
    sigE_ = float(sigE0); sigF_ = float(sigF0); sigH_ = float(sigH0)
    estep = expect_m_dense if debug else expect_m
    logl_incr = np.full(EMsteps, np.nan)
    outp = None
    i = 0
    for i in range(1, EMsteps + 1):                                          # :781
        outp = maximiz_m(estep(X, Y, W_, C_, B_, sigE_, sigF_, sigH_, sigT_), X, Y, type)  # :782
        W_ = outp["W"]; C_ = outp["C"]; B_ = outp["B"]                       # :783-785
        sigE_ = outp["sigE"]; sigF_ = outp["sigF"]; sigH_ = outp["sigH"]; sigT_ = outp["sigT"]
        logl_incr[i - 1] = logl_w(X, Y, W_, C_, B_, sigE_, sigF_, sigH_, sigT_)   # :791
        if record is not None:
            record(i, dict(W=W_, C=C_, B=B_, sigE=sigE_, sigF=sigF_, sigH=sigH_, sigT=sigT_,
                           loglik=logl_incr[i - 1]))
        if i > 1 and (logl_incr[i - 1] - logl_incr[i - 2]) < atol:           # :792 (signed)
            break
    Wn, Cn, Bn, Tn = normalise(W_, C_, B_, sigT_)                            # :794-799
    outp = dict(outp)
    outp["W"] = Wn; outp["C"] = Cn; outp["B"] = Bn; outp["sigT"] = Tn
    logl_incr = logl_incr[:i]                                                # :800
    negative_increments = bool(np.any(np.diff(logl_incr) < 0))               # :801 (warning)
    Eout = expect_m(X, Y, W_, C_, B_, sigE_, sigF_, sigH_, sigT_)            # :802 (un-reordered)
    return dict(Expectations=Eout, loglik=logl_incr, estimates=outp,
                loop_state=dict(W=W_, C=C_, B=B_, sigE=sigE_, sigF=sigF_, sigH=sigH_, sigT=sigT_),
                negative_increments=negative_increments)
