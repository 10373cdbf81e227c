"""Host-side mirror of the reference's R interface for the PPLS_simult path, over the C ABI.

R is not installed on the build or GPU boxes, so this module plays the role of the R
wrapper + Rcpp shim (r/PPLS_simult_b200.R, r/ppls_b200_shim.cpp): same
function name, argument meaning, defaults, returned list structure and warnings as
Package/PPLS/R/EM_W_multi.R:758-807.  All arithmetic on X and Y happens in libppls_b200.so
on the GPU; only r-sized bookkeeping (the sign/order normalisation, :773-778 and :794-799)
is done here, exactly as the C ABI header assigns it to the caller.
"""
from __future__ import annotations

import ctypes as C
import warnings
from typing import Optional

import numpy as np

from . import _lib as L

__all__ = ["Context", "Data", "Fit", "PPLS_simult", "PPLS_simult_batched", "PPLS", "fconstraint", "logl_W", "orth", "normalise", "default_init",
           "sequential_init", "default_context", "variances", "simulC", "meta_PPLSi"]


def _f64(a, order="F"):
    return np.require(np.asarray(a, dtype=np.float64), dtype=np.float64, requirements=[order, "A"])


def _diagvec(M, r):
    M = np.asarray(M, dtype=np.float64)
    v = np.diag(M).copy() if M.ndim == 2 else np.atleast_1d(M).astype(np.float64).copy()
    if v.shape != (r,):
        raise ValueError(f"expected {r} diagonal entries, got shape {v.shape}")
    return v


class _ParamsBuf:
    """Owns numpy storage for a ppls_params struct."""

    def __init__(self, p, q, r, W=None, C_=None, B=None, sigT=None, sigE=0.0, sigF=0.0, sigH=0.0):
        self.W = _f64(W if W is not None else np.zeros((p, r))).reshape(p, r, order="F")
        self.C = _f64(C_ if C_ is not None else np.zeros((q, r))).reshape(q, r, order="F")
        if self.W.shape != (p, r) or self.C.shape != (q, r):
            raise ValueError("W must be p x r and C must be q x r")
        self.W = np.asfortranarray(self.W.copy()); self.C = np.asfortranarray(self.C.copy())
        self.B = _diagvec(B if B is not None else np.zeros(r), r)
        self.sigT = _diagvec(sigT if sigT is not None else np.zeros(r), r)
        self.struct = L.Params(self.W.ctypes.data_as(L.c_double_p), self.C.ctypes.data_as(L.c_double_p),
                               self.B.ctypes.data_as(L.c_double_p), self.sigT.ctypes.data_as(L.c_double_p),
                               float(sigE), float(sigF), float(sigH))


class Context:
    """ppls_ctx: one GPU (and, for row-sharded runs, one rank of the NCCL communicator)."""

    def __init__(self, device: int = 0, rank: int = 0, world: int = 1, nccl_id: Optional[bytes] = None):
        self._h = C.c_void_p()
        lib = L.lib()
        if world > 1:
            if nccl_id is None or len(nccl_id) < L.PPLS_NCCL_ID_BYTES:
                raise ValueError("a 128-byte NCCL id from Context.nccl_unique_id() is required for world > 1")
            buf = C.create_string_buffer(bytes(nccl_id), L.PPLS_NCCL_ID_BYTES)
            L.check(lib.ppls_ctx_create_dist(C.byref(self._h), device, rank, world, buf, L.PPLS_NCCL_ID_BYTES))
        else:
            L.check(lib.ppls_ctx_create(C.byref(self._h), device))
        self.device, self.rank, self.world = device, rank, world
        sm = C.c_int()
        L.check(lib.ppls_ctx_info(self._h, None, None, None, C.byref(sm)))
        self.sm_count = sm.value

    @classmethod
    def multi(cls, devices) -> "Context":
        """ONE process driving several GPUs (ppls_ctx_create_multi): rows of X and Y are split over `devices`, the moment
        vector travels by direct peer stores; what `device = 0:7` means in the R wrapper (r/PPLS_simult_b200.R)."""
        devs = [int(d) for d in devices]
        self = cls.__new__(cls)
        self._h = C.c_void_p()
        arr = (C.c_int * len(devs))(*devs)
        L.check(L.lib().ppls_ctx_create_multi(C.byref(self._h), arr, len(devs)))
        self.device, self.rank, self.world, self.devices = devs[0], 0, len(devs), devs
        sm = C.c_int()
        L.check(L.lib().ppls_ctx_info(self._h, None, None, None, C.byref(sm)))
        self.sm_count = sm.value
        return self

    @staticmethod
    def nccl_unique_id() -> bytes:
        buf = C.create_string_buffer(L.PPLS_NCCL_ID_BYTES)
        L.check(L.lib().ppls_nccl_unique_id(buf, L.PPLS_NCCL_ID_BYTES))
        return buf.raw

    def sync(self):
        L.check(L.lib().ppls_ctx_sync(self._h))

    def probe(self, which: str, data: "Optional[Data]" = None) -> float:
        """Measured ceiling of this GPU (ppls_probe): 'hbm_read' -> GB/s of a read-only stream (over `data` when given),
        'fp64_dmma' -> TFLOP/s of the FP64 tensor path.  For roofline reporting only."""
        v = C.c_double(0.0)
        L.check(L.lib().ppls_probe(self._h, data._h if data is not None else None, {"hbm_read": 0, "fp64_dmma": 1}[which], C.byref(v)))
        return float(v.value)

    def close(self):
        if self._h:
            L.lib().ppls_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_default_ctx: Optional[Context] = None


def default_context() -> Context:
    global _default_ctx
    if _default_ctx is None:
        _default_ctx = Context(0)
    return _default_ctx


class Data:
    """ppls_data: a device-resident row shard of [X Y]."""

    def __init__(self, ctx: Context, handle):
        self.ctx, self._h = ctx, handle
        n, p, q, nt = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64()
        sx, sy = C.c_double(), C.c_double()
        L.check(L.lib().ppls_data_info(self._h, C.byref(n), C.byref(p), C.byref(q), C.byref(nt), C.byref(sx), C.byref(sy)))
        self.n, self.p, self.q, self.n_total = n.value, p.value, q.value, nt.value
        self.ssqX, self.ssqY = sx.value, sy.value

    @classmethod
    def from_host(cls, ctx: Context, X, Y) -> "Data":
        X = np.asarray(X); Y = np.asarray(Y)
        if X.ndim != 2 or Y.ndim != 2 or X.shape[0] != Y.shape[0]:
            raise ValueError("X and Y must be matrices with the same number of rows")
        Xf, Yf = _f64(X), _f64(Y)
        n, p = Xf.shape; q = Yf.shape[1]
        h = C.c_void_p()
        L.check(L.lib().ppls_data_from_host(ctx._h, Xf.ctypes.data, max(n, 1), Yf.ctypes.data, max(n, 1), n, p, q, C.byref(h)))
        return cls(ctx, h)

    @classmethod
    def synthetic(cls, ctx: Context, n: int, p: int, q: int, truth: dict, seed: int, row0: int = 0) -> "Data":
        r = np.asarray(truth["W"]).shape[1]
        pb = _ParamsBuf(p, q, r, truth["W"], truth["C"], truth["B"], truth["sigT"], truth["sigE"], truth["sigF"], truth["sigH"])
        h = C.c_void_p()
        L.check(L.lib().ppls_data_synthetic(ctx._h, n, row0, p, q, r, seed, C.byref(pb.struct), C.byref(h)))
        return cls(ctx, h)

    @classmethod
    def simulc(cls, ctx: Context, n: int, truth: dict, P_Yosc=None, P_Xosc=None, sigTo=None, sigUo=None, seed: int = 1,
               row0: int = 0) -> "Data":
        """Resident sample of simulC's model (loglC.cpp:269-315) with optional orthogonal parts."""
        W = _f64(truth["W"]); Cm = _f64(truth["C"])
        p, r = W.shape; q = Cm.shape[0]
        PY = np.asfortranarray(_f64(P_Yosc).reshape(p, -1)) if P_Yosc is not None else np.zeros((p, 0), order="F")
        PX = np.asfortranarray(_f64(P_Xosc).reshape(q, -1)) if P_Xosc is not None else np.zeros((q, 0), order="F")
        nx, ny = PY.shape[1], PX.shape[1]
        to = _diagvec(sigTo if sigTo is not None else np.zeros(nx), nx)
        uo = _diagvec(sigUo if sigUo is not None else np.zeros(ny), ny)
        pb = _ParamsBuf(p, q, r, W, Cm, truth["B"], truth["sigT"], truth["sigE"], truth["sigF"], truth["sigH"])
        h = C.c_void_p()
        L.check(L.lib().ppls_data_simulc(ctx._h, n, row0, p, q, r, nx, ny, seed, C.byref(pb.struct),
                                         PY.ctypes.data if nx else None, PX.ctypes.data if ny else None,
                                         to.ctypes.data if nx else None, uo.ctypes.data if ny else None, C.byref(h)))
        return cls(ctx, h)

    @classmethod
    def from_host_cols(cls, ctx: Context, X_block, Y_block) -> "Data":
        """Column-sharded: this rank's column block of X and of Y (all rows); wide shapes, p >> N."""
        Xf, Yf = _f64(np.asarray(X_block)), _f64(np.asarray(Y_block))
        if Xf.ndim != 2 or Yf.ndim != 2 or Xf.shape[0] != Yf.shape[0]:
            raise ValueError("X and Y blocks must be matrices with the same number of rows")
        n, p = Xf.shape; q = Yf.shape[1]
        h = C.c_void_p()
        L.check(L.lib().ppls_data_from_host_cols(ctx._h, Xf.ctypes.data, max(n, 1), Yf.ctypes.data, max(n, 1), n, p, q, C.byref(h)))
        return cls(ctx, h)

    @classmethod
    def synthetic_cols(cls, ctx: Context, n: int, truth: dict, seed: int, colsX: slice, colsY: slice) -> "Data":
        """Column block [colsX] of X and [colsY] of Y of the synthetic sample (same bits as the row generator)."""
        W = np.asarray(truth["W"])[colsX]; Cm = np.asarray(truth["C"])[colsY]
        r = W.shape[1]
        pb = _ParamsBuf(W.shape[0], Cm.shape[0], r, W, Cm, truth["B"], truth["sigT"], truth["sigE"], truth["sigF"], truth["sigH"])
        h = C.c_void_p()
        L.check(L.lib().ppls_data_synthetic_cols(ctx._h, n, W.shape[0], Cm.shape[0], colsX.start, colsY.start, r, seed,
                                                 C.byref(pb.struct), C.byref(h)))
        return cls(ctx, h)

    def export(self, row0: int = 0, n: Optional[int] = None, out=None):
        """Rows [row0, row0+n) back to the host, column-major.  `out=(X, Y)`: caller-owned column-major
        float64 arrays (e.g. pinned) of shape (n, p) and (n, q) to fill instead of fresh ones."""
        n = self.n - row0 if n is None else n
        if out is not None:
            X, Y = out
            if X.shape != (n, self.p) or Y.shape != (n, self.q) or not (X.flags.f_contiguous and Y.flags.f_contiguous) \
                    or X.dtype != np.float64 or Y.dtype != np.float64:
                raise ValueError("out must be column-major float64 arrays of shape (n, p) and (n, q)")
        else:
            X = np.empty((n, self.p), order="F"); Y = np.empty((n, self.q), order="F")
        L.check(L.lib().ppls_data_export(self.ctx._h, self._h, row0, n, X.ctypes.data, max(n, 1), Y.ctypes.data, max(n, 1)))
        return X, Y

    def close(self):
        if self._h:
            L.lib().ppls_data_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Fit:
    """ppls_fit: one EM run entered at EM_W_multi.R:780 with an explicit start state."""

    def __init__(self, data: Data, init: dict, a: int, EMsteps: int = 10, atol: float = 1e-4, type: str = "SVD",
                 timing: bool = False, check_every: int = 0):
        if type not in ("SVD", "QR"):
            raise ValueError("'arg' should be one of 'SVD', 'QR'")       # match.arg(type), :761
        self.data, self.r, self.EMsteps = data, int(a), int(EMsteps)
        self.opts = L.Opts(int(a), int(EMsteps), L.PPLS_ORTH_SVD if type == "SVD" else L.PPLS_ORTH_QR,
                           int(check_every), float(atol), L.PPLS_F_TIMING if timing else 0, 0)
        pb = _ParamsBuf(data.p, data.q, self.r, init["W"], init["C"], init["B"], init["sigT"],
                        init["sigE"], init["sigF"], init["sigH"])
        self._h = C.c_void_p()
        L.check(L.lib().ppls_fit_create(data.ctx._h, data._h, C.byref(self.opts), C.byref(pb.struct), C.byref(self._h)))

    def iterate(self, n_passes: int):
        done, fin = C.c_int32(), C.c_int32()
        L.check(L.lib().ppls_fit_iterate(self._h, int(n_passes), C.byref(done), C.byref(fin)))
        return done.value, bool(fin.value)

    def run(self):
        fin = False
        while not fin:
            _, fin = self.iterate(self.EMsteps + 1)

    def result(self, want_mu: bool = True) -> dict:
        d, r = self.data, self.r
        est = _ParamsBuf(d.p, d.q, r)
        loglik = np.zeros(max(self.EMsteps, 1))
        steps = C.c_int32()
        mom = np.zeros(3 * r + 2 + r * r)
        muT = np.empty((d.n, r), order="F") if want_mu else None
        muU = np.empty((d.n, r), order="F") if want_mu else None
        L.check(L.lib().ppls_fit_result(self._h, C.byref(est.struct), loglik.ctypes.data, C.byref(steps), mom.ctypes.data,
                                        muT.ctypes.data if want_mu else None, muU.ctypes.data if want_mu else None))
        return dict(W=est.W, C=est.C, B=est.B, sigT=est.sigT, sigE=est.struct.sigE, sigF=est.struct.sigF,
                    sigH=est.struct.sigH, loglik=loglik[:steps.value].copy(), steps=steps.value,
                    Ctt=mom[:r].copy(), Cuu=mom[r:2 * r].copy(), Cut=mom[2 * r:3 * r].copy(), Cee=mom[3 * r], Cff=mom[3 * r + 1],
                    Chh=mom[3 * r + 2:].reshape(r, r, order="F").copy(), mu_T=muT, mu_U=muU)

    def timing(self) -> dict:
        t = L.Timing()
        L.check(L.lib().ppls_fit_timing(self._h, C.byref(t)))
        return {k: getattr(t, k) for k, _ in L.Timing._fields_}

    def close(self):
        if self._h:
            L.lib().ppls_fit_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def normalise(W, C_, B, sigT):
    """Sign/order normalisation of EM_W_multi.R:773-778 (start values) and :794-799 (result),
    quirk included: for W and C the sign vector is applied after the column permutation
    without being permuted; for B it is applied before."""
    W = np.asarray(W, dtype=np.float64); C_ = np.asarray(C_, dtype=np.float64)
    r = W.shape[1]
    b = _diagvec(B, r); t = _diagvec(sigT, r)
    signLoad = np.sign(t * b)
    rotLoad = np.argsort(-(t * b * signLoad), kind="stable")      # order(decreasing=TRUE) is stable
    return (W[:, rotLoad] * signLoad[None, :], C_[:, rotLoad] * signLoad[None, :],
            np.diag((b * signLoad)[rotLoad]), np.diag(t[rotLoad]))


def default_init(p: int, q: int, a: int) -> dict:
    """A data-independent start state (used by tests and benchmarks; PPLS_simult's own default is
    sequential_init): the first `a` cosine basis vectors (orthonormal), B = I,
    sigT = I with a mild decay, sigE = 1/sqrt(p), sigF = 1/sqrt(q), sigH = 0.1."""
    def basis(n):
        j = (np.arange(n) + 0.5)[:, None]
        k = np.arange(1, a + 1)[None, :]
        M = np.cos(np.pi * j * k / n)
        Q, _ = np.linalg.qr(M)
        return Q
    return dict(W=basis(p), C=basis(q), B=np.ones(a), sigT=np.exp(-0.1 * np.arange(a)),
                sigE=1.0 / np.sqrt(p), sigF=1.0 / np.sqrt(q), sigH=0.1)


def sequential_init(data: "Data", a: int, initialGuess: str = "random", seed: int = 0) -> dict:
    """Start state of PPLS_simult (EM_W_multi.R:762-771): PPLS(X, Y, a, 20, 1e-4, 'random') -- 20 steps of the sequential
    fit per component from a random start, run on the GPU over the resident data (ppls_sequential_run), with the
    reference's three attempts (:762-764: a failed fit is retried twice with fresh draws; here attempt k uses seed + k).
    The draws follow the reference's distributions (:133) on the counter-based generator, not R's RNG stream: the same
    control flow and the same family of starts, repeatable for a given seed."""
    last = None
    for attempt in range(3):
        try:
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")                      # suppressWarnings(...), :762
                f0 = PPLS(None, None, a, 20, 1e-4, initialGuess, data=data, loglikelihoods=False, seed=seed + attempt)
            if f0["W"].shape[1] < a:
                raise RuntimeError("sequential start fit stopped early (rank-deficient residuals)")
            sg = f0["sig"]
            return dict(W=f0["W"], C=f0["C"], B=f0["B"], sigT=sg[:, 3], sigE=sg[a - 1, 0], sigF=sg[a - 1, 1], sigH=sg[a - 1, 2])
        except (L.PPLSError, RuntimeError) as ex:
            last = ex
    raise RuntimeError(f"the sequential start fit failed three times ({last}); pass init=")


def PPLS_simult(X, Y, a, EMsteps=10, atol=1e-4, type="SVD", init: Optional[dict] = None,
                ctx: Optional[Context] = None, data: Optional[Data] = None, want_mu: bool = True, seed: int = 0,
                start: str = "random", **kwargs):
    """Simultaneous PPLS fit -- drop-in for PPLS_simult (Package/PPLS/R/EM_W_multi.R:758-807).

    Returns {'Expectations': {mu_T, mu_U, Ctt, Cuu, Cut, Cee, Cff, Chh}, 'loglik', 'estimates':
    {W, C, B, sigE, sigF, sigH, sigT}} with the reference's shapes (B, sigT, Ctt, Cuu, Cut
    r x r diagonal matrices; Cee, Cff 1 x 1).  `init` replaces the start of :762-771 (default: sequential_init, the
    reference's PPLS(X, Y, a, 20, 1e-4, 'random') with its three attempts, drawn from `seed`; start='equal' for a
    deterministic data-independent start).
    `debug=` (the dense Expect_M, :643-667) is accepted and ignored with a warning: the dense
    form is O((p+q)^3) and exists in this repo only inside the CPU oracle."""
    if isinstance(type, (tuple, list)):
        type = type[0]                                   # c("SVD","QR") default -> "SVD"
    if kwargs.pop("debug", False):
        warnings.warn("debug=TRUE (dense Expect_M) is not part of the GPU path; using the fast E-step")
    if kwargs:
        raise TypeError(f"unused arguments: {sorted(kwargs)}")
    a = int(a)
    if a <= 0:
        raise ValueError("#components must be >0")
    if int(EMsteps) < 1:
        raise ValueError("EMsteps must be a strictly positive integer")
    ctx = ctx or (data.ctx if data is not None else default_context())
    own_data = data is None
    if data is None:
        data = Data.from_host(ctx, X, Y)
    try:
        if init is None:
            init = sequential_init(data, a, start, seed)                                   # :762-771
        W0, C0, B0, T0 = normalise(init["W"], init["C"], init["B"], init["sigT"])          # :773-778
        start = dict(W=W0, C=C0, B=B0, sigT=T0, sigE=init["sigE"], sigF=init["sigF"], sigH=init["sigH"])
        fit = Fit(data, start, a, EMsteps, atol, type)
        try:
            fit.run()                                                                      # :780-793
            res = fit.result(want_mu=want_mu)                                              # :802
        finally:
            fit.close()
    finally:
        if own_data:
            data.close()
    Wn, Cn, Bn, Tn = normalise(res["W"], res["C"], res["B"], res["sigT"])                  # :794-799
    loglik = res["loglik"]
    if np.any(np.diff(loglik) < 0):
        warnings.warn("Negative increments of likelihood")                                 # :801
    I = np.eye(a)
    Eout = dict(mu_T=res["mu_T"], mu_U=res["mu_U"], Ctt=np.diag(res["Ctt"]), Cuu=np.diag(res["Cuu"]),
                Cut=np.diag(res["Cut"]), Cee=np.array([[res["Cee"]]]), Cff=np.array([[res["Cff"]]]), Chh=res["Chh"])
    outp = dict(W=Wn, C=Cn, B=Bn, sigE=res["sigE"], sigF=res["sigF"], sigH=res["sigH"], sigT=Tn)
    return dict(Expectations=Eout, loglik=loglik, estimates=outp,
                loop_state=dict(W=res["W"], C=res["C"], B=np.diag(res["B"]), sigT=np.diag(res["sigT"]),
                                sigE=res["sigE"], sigF=res["sigF"], sigH=res["sigH"]),
                **{"class": "PPLS_simult"})


def PPLS_simult_batched(X, Y, a, inits, EMsteps=10, atol=1e-4, type="SVD", ctx: Optional[Context] = None,
                        data: Optional[Data] = None, timing: bool = False):
    """Multi-start PPLS_simult: every element of `inits` is a start state (as `init=` of PPLS_simult);
    all starts run the loop of EM_W_multi.R:780-793 concurrently over ONE resident copy of X and Y
    (ppls_simult_run_batched).  Returns a list of {'loglik', 'estimates', 'loop_state', 'steps'} in the
    order of `inits` (estimates normalised as at :794-799); with timing=True also the batch timing dict.
    The reference has no such entry point: it would call PPLS_simult once per start."""
    if isinstance(type, (tuple, list)):
        type = type[0]
    if type not in ("SVD", "QR"):
        raise ValueError("'arg' should be one of 'SVD', 'QR'")
    a = int(a)
    ctx = ctx or (data.ctx if data is not None else default_context())
    own = data is None
    if data is None:
        data = Data.from_host(ctx, X, Y)
    try:
        S = len(inits)
        starts, outs = [], []
        for init in inits:
            W0, C0, B0, T0 = normalise(init["W"], init["C"], init["B"], init["sigT"])          # :773-778
            starts.append(_ParamsBuf(data.p, data.q, a, W0, C0, B0, T0, init["sigE"], init["sigF"], init["sigH"]))
            outs.append(_ParamsBuf(data.p, data.q, a))
        ini_arr = (L.Params * S)(*[s.struct for s in starts])
        est_arr = (L.Params * S)(*[o.struct for o in outs])
        opts = L.Opts(a, int(EMsteps), L.PPLS_ORTH_SVD if type == "SVD" else L.PPLS_ORTH_QR, 0, float(atol), 0, 0)
        loglik = np.zeros((S, max(int(EMsteps), 1)))
        steps = (C.c_int32 * S)()
        tm = L.Timing()
        L.check(L.lib().ppls_simult_run_batched(ctx._h, data._h, C.byref(opts), S, ini_arr, est_arr,
                                                loglik.ctypes.data, steps, C.byref(tm)))
    finally:
        if own:
            data.close()
    res = []
    for s in range(S):
        o = outs[s]
        Wn, Cn, Bn, Tn = normalise(o.W, o.C, o.B, o.sigT)                                       # :794-799
        e = est_arr[s]
        res.append(dict(loglik=loglik[s, :steps[s]].copy(), steps=int(steps[s]),
                        estimates=dict(W=Wn, C=Cn, B=Bn, sigE=e.sigE, sigF=e.sigF, sigH=e.sigH, sigT=Tn),
                        loop_state=dict(W=o.W, C=o.C, B=np.diag(o.B), sigT=np.diag(o.sigT), sigE=e.sigE, sigF=e.sigF,
                                        sigH=e.sigH)))
    if timing:
        return res, {k: getattr(tm, k) for k, _ in L.Timing._fields_}
    return res


def _guess(p, q, kind, custom, ctx=None, seed=0, component=0):
    """Start state of one component (EM_W_multi.R:126-141): 'equal', 'random' (drawn on the device from the counter-based
    generator keyed by (seed, component): ppls_initial_guess) or 'custom'."""
    if kind == "custom" or custom is not None:
        if custom is None:
            raise ValueError("initialGuess='custom' needs customGuess")
        g = custom
        return _ParamsBuf(p, q, 1, np.reshape(g["W"], (p, 1)), np.reshape(g["C"], (q, 1)), np.ravel(g["B"])[:1],
                          np.ravel(g["sigT"])[:1], g["sigE"], g["sigF"], g["sigH"])
    if kind not in ("equal", "random"):
        raise ValueError("initialGuess must be 'equal', 'random' or 'custom' here: 'o2m' needs OmicsPLS::o2m "
                         "(pass its result as customGuess)")
    g = _ParamsBuf(p, q, 1)
    L.check(L.lib().ppls_initial_guess((ctx or default_context())._h, 0 if kind == "equal" else 1, p, q, int(seed), int(component),
                                       C.byref(g.struct)))
    return g


def fconstraint(constraints=None) -> dict:
    """fconstraint (EM_W_multi.R:85-92): a dict with the keys W, C, B, sigE, sigF, sigH, sigT, each None or the fixed value."""
    outp = dict(W=None, C=None, B=None, sigE=None, sigF=None, sigH=None, sigT=None)
    for k, v in (constraints or {}).items():
        if k in outp:
            outp[k] = v
    return outp


def _constraints_array(constraints, a, p, q):
    """list of nr_comp fconstraint dicts -> (ctypes array of ppls_constraints, keep-alive list)."""
    if constraints is None:
        return None, []
    if len(constraints) != a:
        raise ValueError("There should be a list of constraints for each component, see ?PPLS.")      # :240
    arr = (L.Constraints * a)()
    keep = []
    for i, c in enumerate(constraints):
        c = fconstraint(c)
        mask = 0
        for name, bit in L.FIX_BITS.items():
            v = c[name]
            if v is None:
                continue
            mask |= bit
            if name in ("W", "C"):
                vec = np.ascontiguousarray(np.ravel(v), dtype=np.float64)
                if vec.size != (p if name == "W" else q):
                    raise ValueError(f"constraint on {name}: wrong length")
                keep.append(vec)
                setattr(arr[i], name, vec.ctypes.data_as(L.c_double_p))
            else:
                setattr(arr[i], name, float(np.ravel(v)[0]))
        arr[i].mask = mask
    return arr, keep


def PPLS(X, Y, nr_comp=1, EMsteps=100, atol=1e-4, initialGuess="equal", customGuess=None,
         ctx: Optional[Context] = None, data: Optional[Data] = None, loglikelihoods: bool = True, seed: int = 0,
         critfunc=None, constraints=None):
    """Sequential PPLS fit -- mirror of PPLS (Package/PPLS/R/EM_W_multi.R:229-279) on the GPU: the fit
    PPLS_simult runs for its start values (:762).  critfunc = identity, no constraints.

    Returns {'W' p x a, 'C' q x a, 'B' [a], 'sig' a x 4 (sigX, sigY, sigH, sigT), 'Other_output':
    {'Last_increment', 'Number_steps', 'Loglikelihoods', 'logvalues'}}.  customGuess: one dict (used for
    every component, as the reference does) or a list with one dict per component.  If a noise s.d.
    vanishes at component i the earlier components are returned with a warning (:257-262).  initialGuess='random'
    (:133) draws every component's start on the device from (seed, component) -- the reference's distribution, not
    R's RNG stream.  critfunc (:172): None / "identity" or "abs" (or the builtin abs); constraints: a list with one
    fconstraint dict per component (:230, :142-146, :160-167)."""
    if critfunc in (None, "identity"):
        crit_abs = 0
    elif critfunc in ("abs", abs, np.abs):
        crit_abs = 1
    else:
        raise ValueError("critfunc must be the identity or abs (a flag crosses the C ABI, not an R closure)")
    if isinstance(initialGuess, (tuple, list)):
        initialGuess = initialGuess[0]
    a = int(nr_comp)
    if a <= 0:
        raise ValueError("#components must be >0")
    ctx = ctx or (data.ctx if data is not None else default_context())
    own = data is None
    if data is None:
        data = Data.from_host(ctx, X, Y)
    try:
        p, q = data.p, data.q
        if p < a or q < a:
            raise ValueError("ncol(X) >= nr_comp and ncol(Y) >= nr_comp are required")
        gs = [_guess(p, q, initialGuess, customGuess[i] if isinstance(customGuess, (list, tuple)) else customGuess,
                     ctx=ctx, seed=seed, component=i) for i in range(a)]
        garr = (L.Params * a)(*[g.struct for g in gs])
        out = _ParamsBuf(p, q, a)
        sig = np.zeros((a, 4), order="F")
        logv = np.zeros((a, int(EMsteps) + 1))
        steps = (C.c_int32 * a)()
        last = np.zeros(a)
        ndone = C.c_int32(0)
        carr, _keep = _constraints_array(constraints, a, p, q)
        L.check(L.lib().ppls_sequential_run_ex(ctx._h, data._h, a, int(EMsteps), float(atol), garr, carr, crit_abs,
                                               C.byref(out.struct), sig.ctypes.data, logv.ctypes.data, steps, last.ctypes.data,
                                               C.byref(ndone)))
        k = int(ndone.value)
        if k < a:
            warnings.warn(f"From component {k + 1} on the residuals are of rank < 1e-14 and calculations are stopped.")
        W, Cm, B, sg = out.W[:, :k].copy(), out.C[:, :k].copy(), out.B[:k].copy(), np.ascontiguousarray(sig[:k])
        other = dict(Last_increment=last[:k].copy(), Number_steps=np.array(steps[:k], dtype=int),
                     logvalues=[logv[i, :steps[i] + 1].copy() for i in range(k)])
        if loglikelihoods:                                                                       # :275
            other["Loglikelihoods"] = np.array([
                logl_W(None, None, W[:, :i + 1], Cm[:, :i + 1], B[:i + 1], sg[i, 0], sg[i, 1], sg[i, 2], sg[:i + 1, 3],
                       data=data) for i in range(k)])
    finally:
        if own:
            data.close()
    return dict(W=W, C=Cm, B=B, sig=sg, Other_output=other, **{"class": "PPLS"})


def logl_W(X, Y, W, C_, B_T, sigX, sigY, sigH, sigT, ctx: Optional[Context] = None, data: Optional[Data] = None):
    """Log-likelihood of a parameter set: logl_W (EM_W_multi.R:296-326) -> loglC_fast (loglC.cpp:318-338),
    one streaming pass on the GPU.  B_T and sigT: r x r diagonal matrices or length-r vectors."""
    W = _f64(W); W = W.reshape(W.shape[0], -1)
    r = W.shape[1]
    ctx = ctx or (data.ctx if data is not None else default_context())
    own = data is None
    if data is None:
        data = Data.from_host(ctx, X, Y)
    try:
        th = _ParamsBuf(data.p, data.q, r, W, np.reshape(_f64(C_), (data.q, r), order="F"), B_T, sigT,
                        float(np.ravel(sigX)[0]), float(np.ravel(sigY)[0]), float(np.ravel(sigH)[0]))
        out = C.c_double(0.0)
        L.check(L.lib().ppls_loglik(ctx._h, data._h, r, C.byref(th.struct), C.byref(out)))
    finally:
        if own:
            data.close()
    return float(out.value)


def orth(G, type="SVD", ctx: Optional[Context] = None):
    """OmicsPLS::orth(G, type) as the M-step calls it (EM_W_multi.R:732-733; in-tree copy Package/functions.R:252-260):
    type "SVD" -> the polar factor U V' of the thin SVD of G; "QR" -> qr.Q(qr(G)) * sign(e[,1]' G[,1])."""
    if type not in ("SVD", "QR"):
        raise ValueError("'arg' should be one of 'SVD', 'QR'")
    Gf = _f64(np.asarray(G))
    if Gf.ndim != 2:
        raise ValueError("G must be a matrix")
    n, r = Gf.shape
    ctx = ctx or default_context()
    out = np.empty((n, r), order="F")
    L.check(L.lib().ppls_orth(ctx._h, Gf.ctypes.data, max(n, 1), n, r, L.PPLS_ORTH_SVD if type == "SVD" else L.PPLS_ORTH_QR,
                              out.ctypes.data, max(n, 1)))
    return out


def variances(fit, data, XorY="X", ctx: Optional[Context] = None, dev_data: Optional[Data] = None, matrices: bool = True):
    """Standard errors of the loadings -- mirror of variances.PPLS_simult(fit, data, XorY)
    (Package/PPLS/R/EM_W_multi.R:830-860) on the GPU (ppls_variances).

    fit: what PPLS_simult returns; data: the X (XorY='X') or Y (XorY='Y') matrix the fit was made on, or None
    with dev_data= a resident Data handle.  Returns a dict with one entry per component 0..a-1 (each
    {'B_exp', 'SSt_exp', 'SSt_star'}, p x p), 'varMatrix' (list of p x p) and 'seLoad' (p x a), as the
    reference's list; matrices=False skips everything but seLoad."""
    if isinstance(XorY, (tuple, list)):
        XorY = XorY[0]
    if XorY not in ("X", "Y"):
        raise ValueError("XorY must be 'X' or 'Y'")
    which = 0 if XorY == "X" else 1
    Ex, est = fit["Expectations"], fit["estimates"]
    mu = np.asfortranarray(_f64(Ex["mu_T" if which == 0 else "mu_U"]))
    n, r = mu.shape
    cd = _diagvec(Ex["Ctt" if which == 0 else "Cuu"], r)
    ctx = ctx or (dev_data.ctx if dev_data is not None else default_context())
    own = dev_data is None
    if dev_data is None:
        M = _f64(data)
        if M.shape[0] != n:
            raise ValueError("data and fit$Expectations have different numbers of rows")
        # only one side is needed: the other is a one-column placeholder
        z = np.zeros((n, 1))
        dev_data = Data.from_host(ctx, M, z) if which == 0 else Data.from_host(ctx, z, M)
    try:
        w = dev_data.p if which == 0 else dev_data.q
        se = np.zeros((w, r), order="F")
        var = np.zeros((r, w, w)) if matrices else None
        e1 = np.zeros((r, w, w)) if matrices else None
        e2 = np.zeros((r, w, w)) if matrices else None
        bexp = np.zeros(r)
        Wp = np.zeros((w, r), order="F")
        ptr = lambda a: a.ctypes.data if a is not None else None     # noqa: E731
        L.check(L.lib().ppls_variances(ctx._h, dev_data._h, which, r, mu.ctypes.data, mu.shape[0], cd.ctypes.data,
                                       float(np.ravel(est["sigE"])[0]), se.ctypes.data, ptr(var), ptr(e1), ptr(e2),
                                       bexp.ctypes.data, Wp.ctypes.data))
    finally:
        if own:
            dev_data.close()
    out = dict(seLoad=se, W=Wp)
    if matrices:
        # blocks arrive column-major; every one of these matrices is symmetric, .T restores the orientation
        for i in range(r):
            out[i] = dict(B_exp=bexp[i] * np.eye(w), SSt_exp=e1[i].T, SSt_star=e2[i].T)
        out["varMatrix"] = [var[i].T for i in range(r)]
    return out


def simulC(N, W, C_, P_Yosc, P_Xosc, B_T, sigX, sigY, sigH, sigT, sigTo, sigUo, seed: int = 1, ctx: Optional[Context] = None):
    """Mirror of simulC (Package/PPLS/src/loglC.cpp:269-315; R/RcppExports.R:12-30): returns Dat, N x (p+q),
    rows = samples, columns = X and Y side by side.  Generated on the GPU (ppls_data_simulc) from a
    counter-based generator -- `seed` replaces R's RNG state, which cannot be reproduced outside R.
    B_T, sigT, sigTo, sigUo: diagonal matrices or vectors.  Use Data.simulc to keep the sample resident."""
    ctx = ctx or default_context()
    W = _f64(W); W = W.reshape(W.shape[0], -1)
    truth = dict(W=W, C=_f64(C_).reshape(-1, W.shape[1]), B=B_T, sigT=sigT, sigE=float(np.ravel(sigX)[0]),
                 sigF=float(np.ravel(sigY)[0]), sigH=float(np.ravel(sigH)[0]))
    d = Data.simulc(ctx, int(N), truth, P_Yosc, P_Xosc, sigTo, sigUo, seed)
    try:
        X, Y = d.export()
    finally:
        d.close()
    return np.hstack([X, Y])


def meta_PPLSi(X, Y, Ipopu, EMsteps=100, atol=1e-4, initialGuess="equal", customGuess=None,
               ctx: Optional[Context] = None, data: Optional[Data] = None, pop_rows=None):
    """Multi-population one-component fit -- mirror of meta_PPLSi (Package/PPLS/R/EM_W_multi.R:509-589) on the GPU
    (ppls_meta_run).  Ipopu: population label per row; rows must be grouped by population in the order of the sorted
    labels (the reference slices contiguous blocks by table(Ipopu), :455-456,:551-554).  critfunc = identity, no
    constraints.  Returns {'W', 'C', 'params': [{'B_T','sigX','sigY','sigH','sigT'}, ...], 'log': steps x n_pop}.
    Row-sharded over several ranks (one process per GPU): every rank passes ITS rows and `pop_rows`, the number of its
    rows in each population (same populations in the same order on all ranks, 0 where a rank holds none); the pooled
    sums of :481-482 then run over all ranks (one allreduce per pass)."""
    if isinstance(initialGuess, (tuple, list)):
        initialGuess = initialGuess[0]
    if pop_rows is not None:
        counts = np.asarray(pop_rows, dtype=np.int64)
        lab = np.repeat(np.arange(len(counts)), counts)
    else:
        lab = np.asarray(Ipopu)
        levels, counts = np.unique(lab, return_counts=True)
        if not np.array_equal(lab, np.repeat(levels, counts)):
            raise ValueError("rows must be sorted by population (contiguous blocks in the order of the sorted labels)")
    ctx = ctx or (data.ctx if data is not None else default_context())
    own = data is None
    if data is None:
        data = Data.from_host(ctx, X, Y)
    try:
        if data.n != lab.shape[0]:
            raise ValueError("nrow(X) == length(Ipopu) is required")
        J = len(counts)
        g = _guess(data.p, data.q, initialGuess, customGuess)
        cnt = np.ascontiguousarray(counts, dtype=np.int64)
        W = np.zeros(data.p); Cm = np.zeros(data.q)
        par = np.zeros((J, 5), order="F")
        logv = np.zeros((int(EMsteps), J))
        steps = C.c_int32(0)
        L.check(L.lib().ppls_meta_run(ctx._h, data._h, J, cnt.ctypes.data, int(EMsteps), float(atol), C.byref(g.struct),
                                      W.ctypes.data, Cm.ctypes.data, par.ctypes.data, logv.ctypes.data, C.byref(steps)))
    finally:
        if own:
            data.close()
    k = int(steps.value)
    params = [dict(B_T=par[j, 0], sigX=par[j, 1], sigY=par[j, 2], sigH=par[j, 3], sigT=par[j, 4]) for j in range(J)]
    return dict(W=W, C=Cm, params=params, log=logv[:k].copy(), Number_steps=k)
