"""ctypes binding of oracle/_ref/libppls_ref.so -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

The library is the reference's own native file (Package/PPLS/src/loglC.cpp) compiled UNMODIFIED against the
stand-in Eigen/Rcpp headers of oracle/ref_shim/ (recipe: `make -C oracle ref`; needs /root/reference, so it is
built in the build container and travels to the GPU box as a prebuilt file).  It PINS the oracle: loglC_fast
(row a6 of SURVEY.md 8a), EMstepC_fast (8f row 1), meta_Estep / meta_Mstep (8f row 4), and -- through the dense
loglC / EMstepC of the same file -- the closed-form log-likelihood and the E-step moments of Expect_M (a2).
Only tests/ may import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "_ref", "libppls_ref.so")
REF_SRC = "/root/reference/Package/PPLS/src/loglC.cpp"

_L = None


def available() -> bool:
    return os.path.exists(SO) or os.path.exists(REF_SRC)


def load():
    """Build (when the reference sources are present and the library is missing or stale) and load."""
    global _L
    if _L is not None:
        return _L
    deps = [os.path.join(HERE, "ref_entry.cpp"), os.path.join(HERE, "ref_shim", "RcppEigen.h"),
            os.path.join(HERE, "ref_shim", "Eigen", "Dense")]
    if os.path.exists(REF_SRC) and (not os.path.exists(SO) or any(os.path.getmtime(d) > os.path.getmtime(SO) for d in deps)):
        subprocess.check_call(["make", "-C", HERE, "ref"])
    if not os.path.exists(SO):
        raise RuntimeError("oracle/_ref/libppls_ref.so is missing and /root/reference is not present to build it")
    L = C.CDLL(SO)
    L.ppls_ref_error.restype = C.c_char_p
    L.ppls_ref_source.restype = C.c_char_p
    L.ppls_ref_seed.argtypes = [C.c_uint64]
    L.ppls_ref_last_len.restype = C.c_long
    L.ppls_ref_last_len.argtypes = [C.c_char_p]
    L.ppls_ref_last_get.argtypes = [C.c_char_p, C.c_void_p]
    vp, lg, db = C.c_void_p, C.c_long, C.c_double
    L.ppls_ref_loglC_fast.argtypes = [vp, vp, vp, vp, lg, lg, lg, lg, db, db, vp, vp, vp, vp, vp, vp]
    L.ppls_ref_EMstepC_fast.argtypes = [vp, vp, db, vp, vp, lg, lg, lg, db, db, db, db, db, db, db]
    L.ppls_ref_meta_step.argtypes = [vp, vp, db, vp, vp, lg, lg, lg, db, db, db, db, db, db, db, C.c_int]
    L.ppls_ref_loglC.argtypes = [vp, vp, vp, vp, vp, vp, lg, lg, lg, lg, lg, lg, db, db, db, vp, vp, vp, db, db, db]
    L.ppls_ref_EMstepC.argtypes = [vp, vp, vp, vp, vp, vp, lg, lg, lg, lg, lg, lg, db, db, db, vp, vp, vp]
    L.ppls_ref_simulC.argtypes = [lg, vp, vp, vp, vp, vp, lg, lg, lg, lg, lg, db, db, db, vp, vp, vp, vp]
    _L = L
    return L


def _f(a):
    return np.asfortranarray(np.asarray(a, dtype=np.float64))


def _check(L, rc):
    if rc != 0:
        raise RuntimeError("reference native code: " + L.ppls_ref_error().decode())


def _last(L, key, shape=None):
    n = L.ppls_ref_last_len(key.encode())
    if n < 0:
        raise KeyError(key)
    out = np.zeros(max(n, 1))
    _check(L, L.ppls_ref_last_get(key.encode(), out.ctypes.data))
    out = out[:n]
    if shape is not None:
        return out.reshape(shape, order="F")
    return float(out[0]) if n == 1 else out


def _diagm(v):
    v = np.asarray(v, dtype=np.float64)
    return _f(np.diag(np.atleast_1d(v)) if v.ndim < 2 else v)


def loglC_fast(W, Cm, X, Y, sigX, sigY, sig2T, c1, c2, c3, Kc):
    """loglC_fast(W, C, X, Y, sigX, sigY, sig2T, c1, c2, c3, Kc), loglC.cpp:319-338."""
    L = load()
    W = _f(W); Cm = _f(Cm); X = _f(X); Y = _f(Y)
    if W.ndim == 1: W = W[:, None]
    if Cm.ndim == 1: Cm = Cm[:, None]
    N, p = X.shape; q = Y.shape[1]; a = W.shape[1]
    v = [np.ascontiguousarray(np.atleast_1d(x), dtype=np.float64) for x in (sig2T, c1, c2, c3, Kc)]
    out = C.c_double(0.0)
    _check(L, L.ppls_ref_loglC_fast(W.ctypes.data, Cm.ctypes.data, X.ctypes.data, Y.ctypes.data, N, p, q, a, float(sigX),
                                    float(sigY), *[x.ctypes.data for x in v], C.byref(out)))
    return float(out.value)


_EM_KEYS = ("B", "Cut", "Ctt", "Cuu", "Cee", "Cff", "Chh")


def EMstepC_fast(W, Cm, B, X, Y, sigX, sigY, sigH, sigT, c1, c2, c3):
    """EMstepC_fast, loglC.cpp:341-397 -> dict with the reference's list names."""
    L = load()
    W = _f(np.ravel(W)); Cm = _f(np.ravel(Cm)); X = _f(X); Y = _f(Y)
    N, p = X.shape; q = Y.shape[1]
    _check(L, L.ppls_ref_EMstepC_fast(W.ctypes.data, Cm.ctypes.data, float(B), X.ctypes.data, Y.ctypes.data, N, p, q,
                                      float(sigX), float(sigY), float(sigH), float(sigT), float(c1), float(c2), float(c3)))
    out = {k: _last(L, k) for k in _EM_KEYS}
    for k in ("mu_T", "mu_U", "W", "C", "sighat", "siglathat"):
        out[k] = np.atleast_1d(_last(L, k))
    return out


def meta_step(W, Cm, B, X, Y, sigX, sigY, sigH, sigT, c1, c2, c3, mstep=False):
    """meta_Estep (loglC.cpp:400-449); with mstep=True meta_Mstep (:452-474) applied to its result."""
    L = load()
    W = _f(np.ravel(W)); Cm = _f(np.ravel(Cm)); X = _f(X); Y = _f(Y)
    N, p = X.shape; q = Y.shape[1]
    _check(L, L.ppls_ref_meta_step(W.ctypes.data, Cm.ctypes.data, float(B), X.ctypes.data, Y.ctypes.data, N, p, q,
                                   float(sigX), float(sigY), float(sigH), float(sigT), float(c1), float(c2), float(c3),
                                   1 if mstep else 0))
    if mstep:
        return dict(B=_last(L, "B"), sighat=_last(L, "sighat"), siglathat=_last(L, "siglathat"),
                    Cxt=np.atleast_1d(_last(L, "Cxt")), Cyu=np.atleast_1d(_last(L, "Cyu")))
    out = {k: _last(L, k) for k in ("Vt", "Vu", "Cut", "Ctt", "Cee", "Cff", "Chh")}
    for k in ("Cxt", "Cyu", "mu_T", "mu_U"):
        out[k] = np.atleast_1d(_last(L, k))
    return out


def _dense_args(W, Cm, P_Yosc, P_Xosc, B_T, sigT, sigTo, sigUo):
    W = _f(W); Cm = _f(Cm)
    p, a = W.shape; q = Cm.shape[0]
    PY = _f(np.reshape(P_Yosc, (p, -1))); PX = _f(np.reshape(P_Xosc, (q, -1)))
    return W, Cm, PY, PX, _diagm(B_T), _diagm(sigT), _diagm(sigTo), _diagm(sigUo), p, q, a, PY.shape[1], PX.shape[1]


def loglC(W, Cm, P_Yosc, P_Xosc, B_T, Dat, sigX, sigY, sigH, sigT, sigTo, sigUo, c1=0.0, c2=0.0, c3=0.0):
    """loglC, loglC.cpp:37-113 -> {'value', 'invS'} (dense covariance + Cholesky when c1 = c2 = c3 = 0)."""
    L = load()
    W, Cm, PY, PX, Bm, Tm, Tom, Uom, p, q, a, nx, ny = _dense_args(W, Cm, P_Yosc, P_Xosc, B_T, sigT, sigTo, sigUo)
    D = _f(Dat); N = D.shape[0]
    _check(L, L.ppls_ref_loglC(W.ctypes.data, Cm.ctypes.data, PY.ctypes.data, PX.ctypes.data, Bm.ctypes.data, D.ctypes.data,
                               N, p, q, a, nx, ny, float(sigX), float(sigY), float(sigH), Tm.ctypes.data, Tom.ctypes.data,
                               Uom.ctypes.data, float(c1), float(c2), float(c3)))
    return dict(value=_last(L, "value"), invS=_last(L, "invS", (p + q, p + q)))


def EMstepC(W, Cm, P_Yosc, P_Xosc, B_T, Dat, sigX, sigY, sigH, sigT, sigTo, sigUo):
    """EMstepC, loglC.cpp:117-250: the dense E-step of the PO2PLS model (joint + orthogonal parts)."""
    L = load()
    W, Cm, PY, PX, Bm, Tm, Tom, Uom, p, q, a, nx, ny = _dense_args(W, Cm, P_Yosc, P_Xosc, B_T, sigT, sigTo, sigUo)
    D = _f(Dat); N = D.shape[0]
    _check(L, L.ppls_ref_EMstepC(W.ctypes.data, Cm.ctypes.data, PY.ctypes.data, PX.ctypes.data, Bm.ctypes.data, D.ctypes.data,
                                 N, p, q, a, nx, ny, float(sigX), float(sigY), float(sigH), Tm.ctypes.data, Tom.ctypes.data,
                                 Uom.ctypes.data))
    return dict(Cxt=_last(L, "Cxt", (p, a)), Cyu=_last(L, "Cyu", (q, a)), Ctt=_last(L, "Ctt", (a, a)), Cuu=_last(L, "Cuu", (a, a)),
                Cut=_last(L, "Cut", (a, a)), Chh=_last(L, "Chh", (a, a)), diag_Cee=np.atleast_1d(_last(L, "diag_Cee")),
                diag_Cff=np.atleast_1d(_last(L, "diag_Cff")))


def simulC(N, W, Cm, P_Yosc, P_Xosc, B_T, sigX, sigY, sigH, sigT, sigTo, sigUo, seed=1):
    """simulC, loglC.cpp:269-315, on the stand-in rnorm (not R's RNG): N x (p+q)."""
    L = load()
    W, Cm, PY, PX, Bm, Tm, Tom, Uom, p, q, a, nx, ny = _dense_args(W, Cm, P_Yosc, P_Xosc, B_T, sigT, sigTo, sigUo)
    L.ppls_ref_seed(int(seed))
    D = np.zeros((int(N), p + q), order="F")
    _check(L, L.ppls_ref_simulC(int(N), W.ctypes.data, Cm.ctypes.data, PY.ctypes.data, PX.ctypes.data, Bm.ctypes.data, p, q, a,
                                nx, ny, float(sigX), float(sigY), float(sigH), Tm.ctypes.data, Tom.ctypes.data, Uom.ctypes.data,
                                D.ctypes.data))
    return D
