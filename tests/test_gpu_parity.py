"""GPU parity tests: the CUDA path, called through the C ABI, against the golden vectors of
the CPU oracle and against the oracle run on fresh seeded inputs."""
import numpy as np
import pytest

from helpers import (TOL_LOGLIK, TOL_PARAM, golden_cases, init_of, load_golden, loglik_tol, moment_tol, relmax)

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    import torch
    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    from ppls_b200 import Context
    c = Context(0)
    yield c
    c.close()


def _run(ctx, X, Y, init, r, steps, atol, typ, want_mu=True):
    from ppls_b200 import Data, Fit
    d = Data.from_host(ctx, X, Y)
    f = Fit(d, init, r, steps, atol, typ)
    f.run()
    res = f.result(want_mu=want_mu)
    f.close(); d.close()
    return res


def _compare(res, g, name):
    assert res["steps"] == len(g["loglik"]), (res["steps"], len(g["loglik"]))
    assert relmax(res["loglik"], g["loglik"]) < loglik_tol(name)
    for key, ref in (("W", "loop_W"), ("C", "loop_C"), ("B", "loop_B"), ("sigT", "loop_sigT")):
        assert relmax(res[key], g[ref]) < TOL_PARAM, key
    # est_sig holds (sigE, sigF, sigH) of the final state
    for key, ref in (("sigE", 0), ("sigF", 1), ("sigH", 2)):
        assert abs(res[key] / float(g["est_sig"][ref]) - 1) < TOL_PARAM, key
    mt = moment_tol(name)
    for key in ("Ctt", "Cuu", "Cut", "Chh", "mu_T", "mu_U"):
        assert relmax(res[key], g[key]) < mt, key
    assert abs(res["Cee"] / float(g["Cee"]) - 1) < mt and abs(res["Cff"] / float(g["Cff"]) - 1) < mt


@pytest.mark.parametrize("name", golden_cases())
def test_golden(ctx, name):
    """Same inputs, same start, same number of EM steps as the oracle: parameters <= 1e-8,
    log-likelihood <= 1e-10 relative (BASELINE.json north_star)."""
    g = load_golden(name)
    if str(g["type"]) == "QR":
        res = _run(ctx, g["X"], g["Y"], init_of(g), int(g["r"]), int(g["EMsteps"]), float(g["atol"]), "QR")
    else:
        res = _run(ctx, g["X"], g["Y"], init_of(g), int(g["r"]), int(g["EMsteps"]), float(g["atol"]), "SVD")
    _compare(res, g, name)


def test_public_api_matches_reference_structure(ctx):
    """PPLS_simult(X, Y, a, EMsteps, atol, type): same list structure and normalised estimates."""
    from ppls_b200 import PPLS_simult
    g = load_golden("c1_wellcond_1000")
    out = PPLS_simult(g["X"], g["Y"], int(g["r"]), EMsteps=1000, atol=-np.inf, type=("SVD", "QR"), init=init_of(g), ctx=ctx)
    assert set(out["Expectations"]) == {"mu_T", "mu_U", "Ctt", "Cuu", "Cut", "Cee", "Cff", "Chh"}
    assert set(out["estimates"]) == {"W", "C", "B", "sigE", "sigF", "sigH", "sigT"}
    assert out["class"] == "PPLS_simult"
    assert relmax(out["estimates"]["W"], g["est_W"]) < TOL_PARAM
    assert relmax(out["estimates"]["C"], g["est_C"]) < TOL_PARAM
    assert relmax(np.diag(out["estimates"]["B"]), g["est_B"]) < TOL_PARAM
    assert relmax(np.diag(out["estimates"]["sigT"]), g["est_sigT"]) < TOL_PARAM
    assert out["estimates"]["B"].shape == (2, 2) and out["Expectations"]["Cee"].shape == (1, 1)
    assert relmax(out["loglik"], g["loglik"]) < 1e-10


def test_synthetic_generator_bit_exact(ctx, oracle_lib):
    """Device generator == host restatement, bit for bit, for any row offset (sharding)."""
    from oracle import cbind
    from ppls_b200 import Data, synth
    truth = synth.truth_params(37, 14, 3)
    d = Data.synthetic(ctx, 1000, 37, 14, truth, seed=99, row0=12345)
    X, Y = d.export()
    Xo, Yo = cbind.gen_data(oracle_lib, 99, 12345, 1000, truth)
    assert np.array_equal(X, Xo) and np.array_equal(Y, Yo)
    sx, sy = float(np.sum(Xo * Xo)), float(np.sum(Yo * Yo))
    assert abs(d.ssqX / sx - 1) < 1e-13 and abs(d.ssqY / sy - 1) < 1e-13
    d.close()


def test_export_roundtrip(ctx):
    from ppls_b200 import Data
    rng = np.random.default_rng(0)
    X = rng.standard_normal((301, 33)); Y = rng.standard_normal((301, 7))
    d = Data.from_host(ctx, X, Y)
    X2, Y2 = d.export()
    assert np.array_equal(X, X2) and np.array_equal(Y, Y2)
    X3, Y3 = d.export(100, 57)
    assert np.array_equal(X[100:157], X3) and np.array_equal(Y[100:157], Y3)
    d.close()


@pytest.mark.parametrize("shape", [(20000, 500, 100, 3), (5003, 1000, 200, 3), (4099, 257, 63, 5), (9000, 64, 64, 4)])
def test_fresh_inputs_vs_fused_oracle(ctx, oracle_lib, shape):
    """Device-generated data at bench-like row widths, a few EM steps, against the fused CPU
    restatement run on the exported data."""
    from oracle import ppls_fused as F
    from ppls_b200 import Data, Fit, synth
    N, p, q, r = shape
    truth = synth.truth_params(p, q, r, 0.5, 0.5, 0.3)
    d = Data.synthetic(ctx, N, p, q, truth, seed=4242)
    st = synth.start_params(truth, 3)
    f = Fit(d, st, r, 6, -np.inf, "SVD")
    f.run()
    res = f.result()
    X, Y = d.export()
    o = F.fused_loop(oracle_lib, X, Y, st["W"], st["C"], st["B"], st["sigE"], st["sigF"], st["sigH"], st["sigT"],
                     EMsteps=6, atol=-np.inf)
    assert relmax(res["loglik"], o["loglik"]) < 1e-10
    for k in ("W", "C", "B", "sigT", "Ctt", "Cuu", "Cut", "Chh", "mu_T", "mu_U"):
        assert relmax(res[k], o[k]) < TOL_PARAM, k
    f.close(); d.close()


def test_error_behaviour(ctx):
    """Argument checks mirror the reference's stopifnot()s (EM_W_multi.R:237-240)."""
    from ppls_b200 import Data, Fit, PPLSError, default_init
    rng = np.random.default_rng(1)
    d = Data.from_host(ctx, rng.standard_normal((50, 6)), rng.standard_normal((50, 4)))
    bad = dict(W=np.zeros((6, 5)), C=np.zeros((4, 5)), B=np.ones(5), sigT=np.ones(5), sigE=1.0, sigF=1.0, sigH=1.0)
    with pytest.raises(PPLSError):
        Fit(d, bad, 5)                                # nr_comp > ncol(Y): stopifnot(ncol(Y) >= nr_comp)
    with pytest.raises(ValueError):
        Fit(d, default_init(6, 4, 2), 2, type="PCA")
    with pytest.raises(ValueError):
        Data.from_host(ctx, np.zeros((5, 3)), np.zeros((6, 2)))
    d.close()


@pytest.mark.parametrize("name", ["c1_wellcond_1000", "r5_odd_shapes", "r3_qr", "r8_max", "r1_p_lt_q", "r3_wide_row"])
@pytest.mark.parametrize("panel,single", [("0", "0"), ("1", "0"), ("1", "1")])
def test_golden_through_the_wide_path(ctx, name, panel, single, monkeypatch):
    """The wide-row path forced on the golden cases: the two-pass form with row-major (em_wide.cu) and panel-major storage
    (em_wide2.cu, FP64 tensor pipe), and the single-pass L2-pipelined form (em_wide3.cu): same bar as the fused kernel."""
    monkeypatch.setenv("PPLS_FORCE_WIDE", "1")
    monkeypatch.setenv("PPLS_PANEL", panel)
    monkeypatch.setenv("PPLS_WIDE_SINGLE", single)
    g = load_golden(name)
    res = _run(ctx, g["X"], g["Y"], init_of(g), int(g["r"]), int(g["EMsteps"]), float(g["atol"]), str(g["type"]))
    _compare(res, g, name)


@pytest.mark.parametrize("single", ["1", "0"])
@pytest.mark.parametrize("shape,steps", [((300, 6000, 500, 5), 6), ((1037, 9001, 33, 3), 4), ((2000, 50000, 1000, 5), 3),
                                         ((5000, 2500, 40, 2), 4), ((129, 80000, 300, 8), 3)])
def test_wide_shapes_vs_fused_oracle(ctx, oracle_lib, shape, steps, single, monkeypatch):
    """Rows too wide for the fused kernel's tile (incl. BASELINE config 4 at full size: N=2000, p=50000,
    q=1000, r=5) take the wide path by themselves -- single-pass form (several row lanes when there are fewer panels than
    SMs, up to 5 panels per CTA, ragged last chunk) and two-pass form; compared with the CPU restatement on the exported data."""
    from oracle import ppls_fused as F
    from ppls_b200 import Data, Fit, synth
    monkeypatch.setenv("PPLS_WIDE_SINGLE", single)
    N, p, q, r = shape
    truth = synth.truth_params(p, q, r, 0.5, 0.5, 0.3)
    d = Data.synthetic(ctx, N, p, q, truth, seed=777)
    st = synth.start_params(truth, 3)
    f = Fit(d, st, r, steps, -np.inf, "SVD")
    f.run()
    res = f.result()
    X, Y = d.export()
    o = F.fused_loop(oracle_lib, X, Y, st["W"], st["C"], st["B"], st["sigE"], st["sigF"], st["sigH"], st["sigT"],
                     EMsteps=steps, atol=-np.inf)
    assert relmax(res["loglik"], o["loglik"]) < 1e-10
    for k in ("W", "C", "B", "sigT", "Ctt", "Cuu", "Cut", "Chh", "mu_T", "mu_U"):
        assert relmax(res[k], o[k]) < TOL_PARAM, k
    f.close(); d.close()


@pytest.mark.parametrize("shape,S,steps", [((3000, 60, 22, 2), 9, 12), ((1500, 40, 30, 3), 5, 8), ((800, 33, 17, 8), 3, 5),
                                            ((2000, 50, 20, 1), 11, 10),
                                            # S*r >= 64: the two-GEMM path (em_batch.cu) -- one group, padded group,
                                            # r not dividing 128, several groups, rows not a multiple of the tile
                                            ((3001, 200, 50, 2), 64, 6), ((1500, 131, 37, 3), 25, 5),
                                            ((999, 60, 22, 2), 70, 4), ((700, 300, 140, 8), 20, 3), ((2000, 50, 20, 1), 80, 5),
                                            # column blocks that straddle the X|Y boundary (p2 % 8 == 0)
                                            ((1200, 72, 40, 2), 40, 4), ((900, 136, 100, 4), 16, 3)])
def test_batched_multistart_matches_single_runs(ctx, oracle_lib, shape, S, steps):
    """ppls_simult_run_batched (starts packed into the streaming kernel's component dimension) ==
    one ordinary run per start, and == the oracle for the first start."""
    from oracle import ppls_fused as F
    from ppls_b200 import Data, Fit, PPLS_simult_batched, synth, normalise
    N, p, q, r = shape
    truth = synth.truth_params(p, q, r, 0.5, 0.5, 0.3)
    d = Data.synthetic(ctx, N, p, q, truth, seed=31)
    inits = [synth.start_params(truth, seed=100 + s, noise=0.5 + 0.1 * s) for s in range(S)]
    out = PPLS_simult_batched(None, None, r, inits, EMsteps=steps, atol=1e-3, data=d)
    assert len(out) == S
    for s in range(S):
        W0, C0, B0, T0 = normalise(inits[s]["W"], inits[s]["C"], inits[s]["B"], inits[s]["sigT"])
        st = dict(inits[s], W=W0, C=C0, B=B0, sigT=T0)
        f = Fit(d, st, r, steps, 1e-3, "SVD"); f.run(); one = f.result(want_mu=False); f.close()
        assert out[s]["steps"] == one["steps"]
        assert relmax(out[s]["loglik"], one["loglik"]) < 1e-12
        ls = out[s]["loop_state"]
        for k in ("W", "C"):
            assert relmax(ls[k], one[k]) < 1e-10, k
        assert relmax(np.diag(ls["B"]), one["B"]) < 1e-10 and abs(ls["sigE"] / one["sigE"] - 1) < 1e-10
    X, Y = d.export()
    W0, C0, B0, T0 = normalise(inits[0]["W"], inits[0]["C"], inits[0]["B"], inits[0]["sigT"])
    o = F.fused_loop(oracle_lib, X, Y, W0, C0, B0, inits[0]["sigE"], inits[0]["sigF"], inits[0]["sigH"], T0,
                     EMsteps=steps, atol=1e-3, want_mu=False)
    assert out[0]["steps"] == o["steps"] and relmax(out[0]["loglik"], o["loglik"]) < 1e-10
    assert relmax(out[0]["loop_state"]["W"], o["W"]) < TOL_PARAM
    d.close()


@pytest.mark.parametrize("shape,S", [((3001, 200, 50, 2), 64), ((1111, 50, 20, 1), 80), ((900, 136, 100, 4), 20), ((700, 300, 140, 8), 20),
                                     # more (row tile, component block) tiles than SMs: every CTA of the persistent kernel walks several
                                     ((20011, 72, 40, 2), 64), ((30000, 40, 24, 4), 40)])
@pytest.mark.parametrize("stages,ws,pp", [("2", "1", "0"), ("3", "1", "0"), ("4", "1", "0"), ("3", "0", "0"), ("3", "1", "1")])
def test_batched_fused_projection_matches_unfused(ctx, shape, S, stages, ws, pp):
    """em_batch.cu: the fused projection + posterior-mean + Gram kernel (r divides 8) against the separate
    bproj/bmu kernels, for every cp.async depth of the accumulate GEMM, warp-specialised (producer warp +
    mbarriers) and __syncthreads forms, and the fused kernel as one 128-row tile per CTA (the default) or as two half-CTAs out of phase
    (pp = 1, an experiment that measured slower): same EM trajectory up to rounding."""
    import os
    from ppls_b200 import Data, PPLS_simult_batched, synth
    N, p, q, r = shape
    truth = synth.truth_params(p, q, r, 0.5, 0.5, 0.3)
    d = Data.synthetic(ctx, N, p, q, truth, seed=77)
    inits = [synth.start_params(truth, seed=300 + s, noise=0.4 + 0.05 * s) for s in range(S)]
    outs = {}
    old = {k: os.environ.get(k) for k in ("PPLS_BATCH_FUSED", "PPLS_BACC_STAGES", "PPLS_BGEMM_WS", "PPLS_BPROJ_PP")}
    try:
        os.environ["PPLS_BPROJ_PP"] = pp
        for fused in ("0", "1"):
            os.environ["PPLS_BATCH_FUSED"] = fused
            os.environ["PPLS_BACC_STAGES"] = stages if fused == "1" else "2"
            os.environ["PPLS_BGEMM_WS"] = ws if fused == "1" else "0"
            outs[fused] = PPLS_simult_batched(None, None, r, inits, EMsteps=5, atol=-np.inf, data=d)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    for a, b in zip(outs["0"], outs["1"]):
        assert a["steps"] == b["steps"]
        assert relmax(b["loglik"], a["loglik"]) < 1e-12
        for k in ("W", "C"):
            assert relmax(b["loop_state"][k], a["loop_state"][k]) < 1e-10, k
        assert relmax(np.diag(b["loop_state"]["B"]), np.diag(a["loop_state"]["B"])) < 1e-10
    d.close()


def test_column_sharded_two_gpus():
    """Column-sharded wide path over 2 ranks (one process per GPU, NCCL inside the library) vs the oracle;
    skipped on a 1-GPU box (tests/mgpu_colshard.py is the torchrun entry)."""
    import os, subprocess, sys, torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                        "127.0.0.1", "--master-port", "29541", os.path.join(root, "tests", "mgpu_colshard.py")],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "COLSHARD OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_row_sharded_two_gpus_peer_exchange():
    """Row-sharded EM over 2 ranks with the one-shot peer exchange (and the NCCL allreduce beside it) vs the
    oracle; skipped on a 1-GPU box (tests/mgpu_rowshard.py is the torchrun entry)."""
    import os, subprocess, sys, torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                        "127.0.0.1", "--master-port", "29543", os.path.join(root, "tests", "mgpu_rowshard.py")],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "ROWSHARD OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_peer_exchange_recovers_from_an_abandoned_fit():
    """A fit that one rank abandons mid-run times out on the other rank (PPLS_E_NCCL), and the next fits on the same
    contexts are bit-identical to the one before the failure (tests/mgpu_recover.py; both forms of the exchange)."""
    import os, subprocess, sys, torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, PPLS_XCHG_TIMEOUT_MS="400")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                        "127.0.0.1", "--master-port", "29549", os.path.join(root, "tests", "mgpu_recover.py")],
                       capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0 and "RECOVER OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_multi_device_context_matches_oracle_and_the_torchrun_form(tmp_path):
    """ONE process driving 2 GPUs (ppls_ctx_create_multi; what `device = 0:1` means in the R wrapper): host X, Y split
    over the devices, the EM loop with the peer exchange, results gathered -- against the oracle, against a
    single-device run, and BIT FOR BIT against the one-process-per-GPU form (tests/mgpu_rowshard.py under torchrun)."""
    import os, subprocess, sys, torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from oracle import cbind, ppls_fused as F
    from ppls_b200 import Context, Data, Fit, PPLS_simult, PPLSError, synth
    N, p, q, r, steps = 20001, 301, 77, 3, 12
    truth = synth.truth_params(p, q, r, 0.5, 0.5, 0.3)
    st = synth.start_params(truth, 3)
    Lo = cbind.load()
    X, Y = cbind.gen_data(Lo, 4242, 0, N, truth)
    o = F.fused_loop(Lo, X, Y, st["W"], st["C"], st["B"], st["sigE"], st["sigF"], st["sigH"], st["sigT"], EMsteps=steps, atol=-np.inf)
    mctx = Context.multi([0, 1])
    assert mctx.world == 2
    try:
        for make in ("host", "synthetic"):
            d = Data.from_host(mctx, X, Y) if make == "host" else Data.synthetic(mctx, N, p, q, truth, seed=4242)
            assert (d.n, d.p, d.q, d.n_total) == (N, p, q, N)
            assert abs(d.ssqX / float(np.sum(X * X)) - 1) < 1e-12
            Xb, Yb = d.export(9000, 2500)                       # a row range across the shard boundary
            assert np.array_equal(Xb, X[9000:11500]) and np.array_equal(Yb, Y[9000:11500])
            f = Fit(d, st, r, steps, -np.inf, "SVD")
            f.run()
            res = f.result(want_mu=True)
            f.close(); d.close()
            assert relmax(res["loglik"], o["loglik"]) < TOL_LOGLIK
            for k in ("W", "C", "B", "sigT"):
                assert relmax(res[k], o[k]) < TOL_PARAM, k
            assert relmax(res["mu_T"], o["mu_T"]) < TOL_PARAM and relmax(res["mu_U"], o["mu_U"]) < TOL_PARAM
            if make == "host":
                first = res
            else:                                               # the device generator makes the same bits as the host one
                assert np.array_equal(res["W"], first["W"]) and np.array_equal(res["loglik"], first["loglik"])
        # the public call on the multi-device context, QR type (update kernel K2 with the exchange, not the fused tail)
        out = PPLS_simult(X, Y, r, EMsteps=5, atol=-np.inf, type="QR", init=st, ctx=mctx, want_mu=False)
        one = PPLS_simult(X, Y, r, EMsteps=5, atol=-np.inf, type="QR", init=st, ctx=Context(0), want_mu=False)
        assert relmax(out["loglik"], one["loglik"]) < 1e-12 and relmax(out["estimates"]["W"], one["estimates"]["W"]) < 1e-9
        # entry points that need a collective the multi-device form does not have say so
        with pytest.raises(PPLSError):
            from ppls_b200 import PPLS
            dd = Data.from_host(mctx, X, Y)
            try:
                PPLS(None, None, 2, 5, data=dd)
            finally:
                dd.close()
    finally:
        mctx.close()
    # bit for bit against the one-process-per-GPU form
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    dump = str(tmp_path / "rowshard.npz")
    rr = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                         "127.0.0.1", "--master-port", "29547", os.path.join(root, "tests", "mgpu_rowshard.py")],
                        capture_output=True, text=True, timeout=900, env=dict(os.environ, PPLS_DUMP=dump))
    assert rr.returncode == 0 and "ROWSHARD OK" in rr.stdout, rr.stdout[-2000:] + rr.stderr[-2000:]
    z = np.load(dump)
    assert np.array_equal(z["W"], first["W"]) and np.array_equal(z["C"], first["C"]) and np.array_equal(z["loglik"], first["loglik"])
    assert np.array_equal(z["B"], first["B"]) and np.array_equal(z["sigT"], first["sigT"])


@pytest.mark.parametrize("n,r", [(37, 1), (500, 2), (1201, 3), (50001, 5), (900, 8)])
@pytest.mark.parametrize("cond", [1.0, 30.0, 1e3, 1e5])
def test_orth_matches_the_oracle(ctx, n, r, cond):
    """ppls_orth == OmicsPLS::orth as restated in oracle/ppls_ref.py (EM_W_multi.R:732-733), both types, for singular
    value spreads from orthonormal to cond(G) = 1e5 (cond(G'G) = 1e10): the Newton-Schulz inverse square root of the
    polar factor (cond(G'G) <= ~1e3) and its Jacobi fallback (beyond).  The polar factor is unique, so no sign freedom.
    The engine forms it as G (G'G)^{-1/2}, so its forward error grows like cond(G)^2 eps (the reference's SVD: cond(G)
    eps) -- stated in DESIGN.md; EM iterates have cond(G) of order 1-30."""
    from oracle import ppls_ref as R
    from ppls_b200 import orth
    rng = np.random.default_rng(7 * n + r)
    Q, _ = np.linalg.qr(rng.standard_normal((n, r)))
    V, _ = np.linalg.qr(rng.standard_normal((r, r)))
    sv = np.geomspace(1.0, 1.0 / cond, r) if r > 1 else np.array([3.0])
    G = (Q * sv) @ V.T * 1e3
    out = orth(G, "SVD", ctx=ctx)
    ref = R.orth(G, "SVD")
    tol = max(1e-12, 50 * cond * cond * 2.2e-16)
    assert np.max(np.abs(out - ref)) < tol, (cond, np.max(np.abs(out - ref)))
    assert np.max(np.abs(out.T @ out - np.eye(r))) < 1e-12
    outq = orth(G, "QR", ctx=ctx)
    refq = R.orth(G, "QR")
    assert np.max(np.abs(outq.T @ outq - np.eye(r))) < 1e-12
    assert np.max(np.abs(outq - refq)) < max(1e-11, 50 * cond * 2.2e-16)


def test_orth_rejects_bad_arguments(ctx):
    from ppls_b200 import PPLSError, orth
    with pytest.raises(ValueError):
        orth(np.eye(3), "LU", ctx=ctx)
    with pytest.raises(PPLSError):
        orth(np.ones((2, 3)), "SVD", ctx=ctx)          # more columns than rows
    with pytest.raises(PPLSError):
        orth(np.ones((20, 9)), "SVD", ctx=ctx)         # nr_comp > 8
