"""simulC (Package/PPLS/src/loglC.cpp:269-315) on the GPU: the device generator with orthogonal parts against
its host restatement (bit for bit) and against the model's covariance.  R's RNG cannot be reproduced here, so
parity with the reference is structural (same model, same argument meaning); PARITY UNPINNED."""
import numpy as np
import pytest


def _model(p=23, q=11, r=2, nx=2, ny=1, seed=0):
    g = np.random.default_rng(seed)
    Q = np.linalg.qr(g.standard_normal((p, r + nx)))[0]
    Z = np.linalg.qr(g.standard_normal((q, r + ny)))[0]
    truth = dict(W=Q[:, :r], C=Z[:, :r], B=np.array([1.3, 0.8])[:r], sigT=np.array([1.5, 1.1])[:r], sigE=0.4, sigF=0.3, sigH=0.25)
    return truth, Q[:, r:], Z[:, r:], np.array([0.9, 0.6])[:nx], np.array([0.7])[:ny]


def test_host_generator_without_orthogonal_parts_is_the_plain_one(oracle_lib):
    from oracle import cbind
    truth, PY, PX, to, uo = _model()
    a = cbind.gen_data(oracle_lib, 5, 100, 300, truth)
    b = cbind.gen_data_orth(oracle_lib, 5, 100, 300, truth, PY[:, :0], PX[:, :0], to[:0], uo[:0])
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


def test_host_generator_covariance(oracle_lib):
    """cov(X) = W SigT^2 W' + P_Yosc SigTo^2 P_Yosc' + sigE^2 I, cov(X, Y) = W SigT^2 B C' (simulC's model)."""
    from oracle import cbind
    truth, PY, PX, to, uo = _model()
    n = 200000
    X, Y = cbind.gen_data_orth(oracle_lib, 11, 0, n, truth, PY, PX, to, uo)
    W, Cm, b, t = truth["W"], truth["C"], truth["B"], truth["sigT"]
    Sx = W @ np.diag(t**2) @ W.T + PY @ np.diag(to**2) @ PY.T + truth["sigE"]**2 * np.eye(W.shape[0])
    Sxy = W @ np.diag(t**2 * b) @ Cm.T
    Sy = Cm @ np.diag(t**2 * b**2 + truth["sigH"]**2) @ Cm.T + PX @ np.diag(uo**2) @ PX.T + truth["sigF"]**2 * np.eye(Cm.shape[0])
    tol = 6 * np.max(np.diag(Sx)) / np.sqrt(n)
    assert np.max(np.abs(X.T @ X / n - Sx)) < tol and np.max(np.abs(X.T @ Y / n - Sxy)) < tol
    assert np.max(np.abs(Y.T @ Y / n - Sy)) < tol


@pytest.mark.gpu
@pytest.mark.parametrize("nx,ny", [(2, 1), (0, 1), (2, 0), (0, 0)])
def test_device_generator_bit_exact(oracle_lib, nx, ny):
    from oracle import cbind
    from ppls_b200 import Data, default_context
    truth, PY, PX, to, uo = _model()
    PY, PX, to, uo = PY[:, :nx], PX[:, :ny], to[:nx], uo[:ny]
    ctx = default_context()
    d = Data.simulc(ctx, 777, truth, PY, PX, to, uo, seed=99, row0=4321)
    X, Y = d.export()
    d.close()
    Xo, Yo = cbind.gen_data_orth(oracle_lib, 99, 4321, 777, truth, PY, PX, to, uo)
    assert np.array_equal(X, Xo) and np.array_equal(Y, Yo)


@pytest.mark.gpu
def test_simulc_mirror_signature(oracle_lib):
    """simulC(N, W, C, P_Yosc, P_Xosc, B_T, sigX, sigY, sigH, sigT, sigTo, sigUo) -> Dat (N x (p+q))."""
    from oracle import cbind
    from ppls_b200 import simulC
    truth, PY, PX, to, uo = _model(r=1, nx=1, ny=1)
    Dat = simulC(500, truth["W"], truth["C"], PY, PX, np.diag(truth["B"]), truth["sigE"], truth["sigF"], truth["sigH"],
                 np.diag(truth["sigT"]), np.diag(to), np.diag(uo), seed=3)
    Xo, Yo = cbind.gen_data_orth(oracle_lib, 3, 0, 500, truth, PY, PX, to, uo)
    assert Dat.shape == (500, 23 + 11) and np.array_equal(Dat, np.hstack([Xo, Yo]))
