"""ctypes declarations for oracle/libppls_oracle.so -- TEST INFRASTRUCTURE, NOT PRODUCT CODE."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "libppls_oracle.so")


def declare(L):
    L.ppls_gen_normal.restype = C.c_double
    L.ppls_gen_normal.argtypes = [C.c_uint64, C.c_uint32, C.c_uint64, C.c_uint32]
    L.ppls_gen_uniform.restype = C.c_double
    L.ppls_gen_uniform.argtypes = [C.c_uint64, C.c_uint32, C.c_uint64, C.c_uint32]
    L.ppls_gen_normals.restype = None
    L.ppls_gen_normals.argtypes = [C.c_uint64, C.c_uint32, C.c_uint64, C.c_int64, C.c_int32, C.c_void_p]
    L.ppls_gen_data.restype = None
    L.ppls_gen_data.argtypes = [C.c_uint64, C.c_uint64, C.c_int64, C.c_int32, C.c_int32, C.c_int32,
                                C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_double,
                                C.c_void_p, C.c_int64, C.c_void_p, C.c_int64]
    L.ppls_gen_data_orth.restype = None
    L.ppls_gen_data_orth.argtypes = [C.c_uint64, C.c_uint64, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_double, C.c_double, C.c_double, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64]
    L.ppls_oracle_moments.restype = None
    L.ppls_oracle_moments.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_int32,
                                      C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.ppls_oracle_num_threads.restype = C.c_int
    L.ppls_oracle_set_threads.restype = None
    L.ppls_oracle_set_threads.argtypes = [C.c_int]
    return L


def load(build: bool = True):
    src = os.path.join(HERE, "ppls_oracle.c")
    if build and (not os.path.exists(SO) or os.path.getmtime(SO) < os.path.getmtime(src)):
        subprocess.check_call(["make", "-C", HERE])
    return declare(C.CDLL(SO))


def _f(a):
    return np.asfortranarray(np.asarray(a, dtype=np.float64))


def gen_normals(L, seed, stream, row0, nrows, ncols):
    out = np.empty((nrows, ncols), order="F")
    L.ppls_gen_normals(seed, stream, row0, nrows, ncols, out.ctypes.data)
    return out


def gen_data(L, seed, row0, nrows, truth):
    W = _f(truth["W"]); Cm = _f(truth["C"])
    p, r = W.shape; q = Cm.shape[0]
    b = np.ascontiguousarray(truth["B"], dtype=np.float64); t = np.ascontiguousarray(truth["sigT"], dtype=np.float64)
    X = np.empty((nrows, p), order="F"); Y = np.empty((nrows, q), order="F")
    L.ppls_gen_data(seed, row0, nrows, p, q, r, W.ctypes.data, Cm.ctypes.data, b.ctypes.data, t.ctypes.data,
                    float(truth["sigE"]), float(truth["sigF"]), float(truth["sigH"]),
                    X.ctypes.data, max(nrows, 1), Y.ctypes.data, max(nrows, 1))
    return X, Y


def gen_data_orth(L, seed, row0, nrows, truth, P_Yosc, P_Xosc, sigTo, sigUo):
    """simulC's model with orthogonal parts (loglC.cpp:288-311) on the counter-based generator."""
    W = _f(truth["W"]); Cm = _f(truth["C"])
    p, r = W.shape; q = Cm.shape[0]
    PY = _f(np.reshape(P_Yosc, (p, -1))); PX = _f(np.reshape(P_Xosc, (q, -1)))
    nx, ny = PY.shape[1], PX.shape[1]
    b = np.ascontiguousarray(truth["B"], dtype=np.float64); t = np.ascontiguousarray(truth["sigT"], dtype=np.float64)
    to = np.ascontiguousarray(np.ravel(sigTo), dtype=np.float64); uo = np.ascontiguousarray(np.ravel(sigUo), dtype=np.float64)
    X = np.empty((nrows, p), order="F"); Y = np.empty((nrows, q), order="F")
    L.ppls_gen_data_orth(seed, row0, nrows, p, q, r, nx, ny, W.ctypes.data, Cm.ctypes.data, PY.ctypes.data, PX.ctypes.data,
                         b.ctypes.data, t.ctypes.data, to.ctypes.data, uo.ctypes.data,
                         float(truth["sigE"]), float(truth["sigF"]), float(truth["sigH"]),
                         X.ctypes.data, max(nrows, 1), Y.ctypes.data, max(nrows, 1))
    return X, Y


def moments(L, X, Y, W, Cm, aT, bT, aU, bU, nranks=1, want_mu=False):
    X = _f(X); Y = _f(Y); W = _f(W); Cm = _f(Cm)
    N, p = X.shape; q = Y.shape[1]; r = W.shape[1]
    co = [np.ascontiguousarray(v, dtype=np.float64) for v in (aT, bT, aU, bU)]
    G1 = np.zeros((p, r), order="F"); G2 = np.zeros((q, r), order="F")
    Gam = np.zeros((2 * r, 2 * r), order="F"); ssq = np.zeros(2)
    muT = np.zeros((N, r), order="F") if want_mu else None
    muU = np.zeros((N, r), order="F") if want_mu else None
    L.ppls_oracle_moments(X.ctypes.data, N, Y.ctypes.data, N, N, p, q, r, W.ctypes.data, Cm.ctypes.data,
                          co[0].ctypes.data, co[1].ctypes.data, co[2].ctypes.data, co[3].ctypes.data, nranks,
                          G1.ctypes.data, G2.ctypes.data, Gam.ctypes.data, ssq.ctypes.data,
                          muT.ctypes.data if want_mu else None, muU.ctypes.data if want_mu else None)
    return dict(G1=G1, G2=G2, Gam=Gam, ssqX=ssq[0], ssqY=ssq[1], mu_T=muT, mu_U=muU)


def random_guess(L, p, q, seed, comp):
    """PPLSi's initialGuess = "random" (EM_W_multi.R:133) on the counter-based generator: the host restatement of
    ppls_b200's device kernel (data.cu::guess_kernel) -- streams 16 (W), 17 (C), 18 (chi-square scalars)."""
    u = np.array([L.ppls_gen_uniform(seed, 16, comp, j) for j in range(p)])
    v = np.array([L.ppls_gen_uniform(seed, 17, comp, j) for j in range(q)])

    def chisq(first, k):
        s = 0.0
        for e in range(k):
            z = L.ppls_gen_normal(seed, 18, comp, first + e)
            s += z * z
        return s
    return dict(W=u / np.sqrt(np.sum(u * u)), C=v / np.sqrt(np.sum(v * v)), B=chisq(0, 1), sigH=chisq(1, 100) / 100.0,
                sigT=chisq(101, 100) / 100.0, sigE=chisq(201, 10) / 100.0, sigF=chisq(211, 10) / 100.0)
