"""world_size-2 (and 3) CPU tests over gloo of the row-sharded N>1 path's host-side logic:
the shard boundaries bench.py uses, the counter-based generator producing the same rows for any
sharding, one sum-allreduce of the moment vector per EM iteration, and the replicated update
staying bit-identical on every rank (so no broadcast is needed).  The per-rank data pass here is
the CPU oracle; the GPU engine performs the same exchange with ncclAllReduce (csrc/capi.cu)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    return port


def _worker(rank, world, port, shape, steps, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), OMP_NUM_THREADS="1")
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import math
    from bench import shard_rows
    from oracle import cbind, ppls_fused as F, ppls_ref as R
    from ppls_b200 import synth
    N, p, q, r = shape
    L = cbind.load(build=False)
    truth = synth.truth_params(p, q, r, 0.5, 0.5, 0.3)
    st = synth.start_params(truth, 5)
    lo, n = shard_rows(N, rank, world)
    X, Y = cbind.gen_data(L, 777, lo, n, truth)                   # this rank's rows only
    W, Cm = st["W"].copy(), st["C"].copy()
    b, t = np.asarray(st["B"], float).copy(), np.asarray(st["sigT"], float).copy()
    sigE, sigF, sigH = st["sigE"], st["sigF"], st["sigH"]

    def allreduce(*arrs):
        flat = torch.from_numpy(np.concatenate([np.ravel(a, order="F") for a in arrs]).copy())
        dist.all_reduce(flat, op=dist.ReduceOp.SUM)
        out, o = [], 0
        for a in arrs:
            out.append(flat[o:o + a.size].numpy().reshape(a.shape, order="F").copy()); o += a.size
        return out

    ssq = np.array([np.sum(X * X), np.sum(Y * Y), float(n)])
    (ssq,) = allreduce(ssq)                                       # once per dataset
    assert int(round(ssq[2])) == N
    trace, k = [], 0
    while True:
        _, aT, bT, aU, bU = F.posterior_coef(b, t, sigE, sigF, sigH)
        m = cbind.moments(L, X, Y, W, Cm, aT, bT, aU, bU)
        G1, G2, Gam = allreduce(m["G1"], m["G2"], m["Gam"])       # ONE exchange per EM iteration
        mom = dict(G1=G1, G2=G2, Gam=Gam, ssqX=ssq[0], ssqY=ssq[1])
        ll = F.loglik_from_moments(mom, N, p, q, b, t, sigE, sigF, sigH)
        if k >= 1:
            trace.append(ll)
        E = F.estep_moments(mom, N, p, q, b, t, sigE, sigF, sigH, W.T @ W, Cm.T @ Cm)
        if k >= steps:
            break
        W = R.orth(G1, "SVD"); Cm = R.orth(G2, "SVD")
        b = E["Cut"] * (1.0 / E["Ctt"]); t = np.sqrt(E["Ctt"])
        sigE = math.sqrt(E["Cee"]); sigF = math.sqrt(E["Cff"])
        sigH = math.sqrt(float(np.sum(np.diag(E["Chh"]))) / r)
        k += 1
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), W=W, C=Cm, B=b, sigT=t, sig=np.array([sigE, sigF, sigH]),
             loglik=np.array(trace), lo=lo, n=n)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_row_sharded_em_matches_single_process(tmp_path, oracle_lib, world):
    from oracle import cbind, ppls_fused as F
    from ppls_b200 import synth
    from helpers import relmax
    shape, steps = (1001, 40, 14, 3), 8
    mp.spawn(_worker, args=(world, _free_port(), shape, steps, str(tmp_path)), nprocs=world, join=True)
    N, p, q, r = shape
    truth = synth.truth_params(p, q, r, 0.5, 0.5, 0.3)
    st = synth.start_params(truth, 5)
    X, Y = cbind.gen_data(oracle_lib, 777, 0, N, truth)
    ref = F.fused_loop(oracle_lib, X, Y, st["W"], st["C"], st["B"], st["sigE"], st["sigF"], st["sigH"], st["sigT"],
                       EMsteps=steps, atol=-np.inf, want_mu=False)
    ranks = [np.load(tmp_path / f"rank{k}.npz") for k in range(world)]
    # shards tile [0, N) exactly
    assert ranks[0]["lo"] == 0 and sum(int(z["n"]) for z in ranks) == N
    for a, b_ in zip(ranks[:-1], ranks[1:]):
        assert int(a["lo"]) + int(a["n"]) == int(b_["lo"])
    # every rank holds bit-identical parameters (replicated update, no broadcast needed)
    for z in ranks[1:]:
        for key in ("W", "C", "B", "sigT", "sig", "loglik"):
            assert np.array_equal(z[key], ranks[0][key]), key
    # and they match the single-process run to rounding (sums are grouped differently)
    z = ranks[0]
    assert relmax(z["loglik"], ref["loglik"]) < 1e-12
    for key in ("W", "C", "B", "sigT"):
        assert relmax(z[key], ref[key]) < 1e-10, key
    assert abs(z["sig"][0] / ref["sigE"] - 1) < 1e-10 and abs(z["sig"][2] / ref["sigH"] - 1) < 1e-10


def test_shard_rows_tiles_the_range():
    from bench import shard_rows
    for N in (1, 7, 1000, 10_000_000):
        for world in (1, 2, 3, 4, 8):
            edges = [shard_rows(N, k, world) for k in range(world)]
            assert edges[0][0] == 0 and sum(n for _, n in edges) == N
            for (lo, n), (lo2, _) in zip(edges[:-1], edges[1:]):
                assert lo + n == lo2


def _col_worker(rank, world, port, shape, steps, out_dir):
    """Column-sharded EM over gloo: rank g holds columns block(g) of X and Y; per iteration one allreduce of
    the N x 2r projection sums and one exchange of the (p+q) x r moment rows (zero-padded sum = all-gather)."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), OMP_NUM_THREADS="1")
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import math
    from oracle import cbind, ppls_fused as F, ppls_ref as R
    from ppls_b200 import synth
    N, p, q, r = shape
    L = cbind.load(build=False)
    truth = synth.truth_params(p, q, r, 0.5, 0.5, 0.3)
    st = synth.start_params(truth, 5)
    X, Y = cbind.gen_data(L, 778, 0, N, truth)
    cx = slice(p * rank // world, p * (rank + 1) // world); cy = slice(q * rank // world, q * (rank + 1) // world)
    Xb, Yb = X[:, cx], Y[:, cy]                                   # this rank only ever touches its block
    del X, Y
    W, Cm = st["W"].copy(), st["C"].copy()
    b, t = np.asarray(st["B"], float).copy(), np.asarray(st["sigT"], float).copy()
    sigE, sigF, sigH = st["sigE"], st["sigF"], st["sigH"]

    def allreduce(a):
        tt = torch.from_numpy(np.ascontiguousarray(a).copy()); dist.all_reduce(tt); return tt.numpy()

    ssq = allreduce(np.array([np.sum(Xb * Xb), np.sum(Yb * Yb)]))
    trace, k = [], 0
    while True:
        _, aT, bT, aU, bU = F.posterior_coef(b, t, sigE, sigF, sigH)
        proj = allreduce(np.concatenate([Xb @ W[cx], Yb @ Cm[cy]], axis=1))      # N x 2r exchange
        Xw, Yc = proj[:, :r], proj[:, r:]
        muT, muU = Xw * aT + Yc * bT, Xw * aU + Yc * bU
        G = np.zeros((p + q, r)); G[cx] = Xb.T @ muT; G[p + cy.start:p + cy.stop] = Yb.T @ muU
        G = allreduce(G)                                                            # (p+q) x r exchange
        Gam = np.block([[Xw.T @ Xw, Xw.T @ Yc], [Yc.T @ Xw, Yc.T @ Yc]])
        mom = dict(G1=G[:p], G2=G[p:], Gam=Gam, ssqX=ssq[0], ssqY=ssq[1])
        ll = F.loglik_from_moments(mom, N, p, q, b, t, sigE, sigF, sigH)
        if k >= 1:
            trace.append(ll)
        E = F.estep_moments(mom, N, p, q, b, t, sigE, sigF, sigH, W.T @ W, Cm.T @ Cm)
        if k >= steps:
            break
        W = R.orth(mom["G1"], "SVD"); Cm = R.orth(mom["G2"], "SVD")
        b = E["Cut"] * (1.0 / E["Ctt"]); t = np.sqrt(E["Ctt"])
        sigE = math.sqrt(E["Cee"]); sigF = math.sqrt(E["Cff"]); sigH = math.sqrt(float(np.sum(np.diag(E["Chh"]))) / r)
        k += 1
    np.savez(os.path.join(out_dir, f"crank{rank}.npz"), W=W, C=Cm, B=b, sigT=t, loglik=np.array(trace))
    dist.barrier()
    dist.destroy_process_group()


def test_column_sharded_em_matches_single_process(tmp_path, oracle_lib):
    from oracle import cbind, ppls_fused as F
    from ppls_b200 import synth
    from helpers import relmax
    shape, steps, world = (120, 301, 45, 3), 6, 2
    mp.spawn(_col_worker, args=(world, _free_port(), shape, steps, str(tmp_path)), nprocs=world, join=True)
    N, p, q, r = shape
    truth = synth.truth_params(p, q, r, 0.5, 0.5, 0.3)
    st = synth.start_params(truth, 5)
    X, Y = cbind.gen_data(oracle_lib, 778, 0, N, truth)
    ref = F.fused_loop(oracle_lib, X, Y, st["W"], st["C"], st["B"], st["sigE"], st["sigF"], st["sigH"], st["sigT"],
                       EMsteps=steps, atol=-np.inf, want_mu=False)
    z0, z1 = (np.load(tmp_path / f"crank{k}.npz") for k in range(world))
    for key in ("W", "C", "B", "sigT", "loglik"):
        assert np.array_equal(z0[key], z1[key]), key               # replicated update: bit-identical ranks
    assert relmax(z0["loglik"], ref["loglik"]) < 1e-12
    for key in ("W", "C", "B", "sigT"):
        assert relmax(z0[key], ref[key]) < 1e-10, key
