"""Row-sharded EM on N GPUs (torchrun): one process per GPU, the moment vector exchanged once per EM
iteration -- by the one-shot peer exchange (peer stores + flags, em_update.cu::xsend_kernel) and, for
comparison, by the NCCL allreduce (PPLS_XCHG=0).  Checked on rank 0 against the CPU oracle on the whole data.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tests/mgpu_rowshard.py [N p q r steps]"""
import os, sys
import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from ppls_b200 import Context, Data, Fit, PPLS, PPLS_simult, PPLS_simult_batched, meta_PPLSi, synth   # noqa: E402


def fit_once(local, rank, world, dev, N, p, q, r, steps, truth, st, xchg):
    os.environ["PPLS_XCHG"] = "1" if xchg else "0"
    idt = torch.zeros(128, dtype=torch.uint8, device=dev)
    if rank == 0:
        idt.copy_(torch.frombuffer(bytearray(Context.nccl_unique_id()), dtype=torch.uint8))
    dist.broadcast(idt, 0)
    ctx = Context(local, rank, world, bytes(idt.cpu().numpy().tobytes()))
    lo, hi = N * rank // world, N * (rank + 1) // world
    d = Data.synthetic(ctx, hi - lo, p, q, truth, seed=4242, row0=lo)
    out = []
    for rep in range(2):                       # two fits on one context: the exchange epoch carries over
        f = Fit(d, st, r, steps, -np.inf, "SVD")
        f.run()
        out.append(f.result(want_mu=False))
        f.close()
    seq = None
    if not xchg:
        # the sequential fit (start-value path of PPLS_simult, NCCL exchange) with several components: the deflation
        # of component i+1 reads the Gram of the pass that ended component i, which the host has over-queued
        seq = PPLS(None, None, 3, 15, 1e-4, "equal", data=d, loglikelihoods=False)
        sim = PPLS_simult(None, None, 2, EMsteps=6, atol=-np.inf, data=d, want_mu=False, start="equal")   # init=None: sequential start
        seq["simult_default_start"] = sim
        # batched multi-start on row shards (both data paths: packs through K1, two-GEMM form): start 0 is `st`
        inits = [st] + [synth.start_params(truth, 50 + j) for j in range(2)]
        seq["batched"] = []
        for gemm in ("0", "1"):
            os.environ["PPLS_BATCH_GEMM"] = gemm
            seq["batched"].append(PPLS_simult_batched(None, None, r, inits, EMsteps=steps, atol=-np.inf, data=d))
        os.environ.pop("PPLS_BATCH_GEMM", None)
        # multi-population fit on row shards: three populations cut so that one spans the ranks' boundary and the last
        # rank holds none of the first (local row counts per population, zeros allowed)
        cuts = [0, int(N * 0.31), int(N * 0.66), N]
        local_counts = [max(0, min(hi, cuts[j + 1]) - max(lo, cuts[j])) for j in range(3)]
        seq["meta"] = meta_PPLSi(None, None, None, 8, -np.inf, data=d, pop_rows=local_counts)
        seq["meta_cuts"] = cuts
    d.close()
    dist.barrier()
    ctx.close()
    dist.barrier()
    return out + [seq]


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    N, p, q, r, steps = (int(a) for a in sys.argv[1:6]) if len(sys.argv) >= 6 else (20001, 301, 77, 3, 12)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    truth = synth.truth_params(p, q, r, 0.5, 0.5, 0.3)
    st = synth.start_params(truth, 3)
    xa, xb, _ = fit_once(local, rank, world, dev, N, p, q, r, steps, truth, st, True)
    na, _, seq = fit_once(local, rank, world, dev, N, p, q, r, steps, truth, st, False)
    # every rank must hold bit-identical parameters (rank-ordered sums on every rank)
    t = torch.from_numpy(np.concatenate([xa["W"].ravel(), xa["C"].ravel(), xa["loglik"]])).to(dev)
    ts = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(ts, t)
    same_ranks = all(torch.equal(ts[0], x) for x in ts)
    ok = True
    if rank == 0:
        from oracle import cbind, ppls_fused as F
        from helpers import relmax
        Lo = cbind.load()
        X, Y = cbind.gen_data(Lo, 4242, 0, N, truth)
        o = F.fused_loop(Lo, X, Y, st["W"], st["C"], st["B"], st["sigE"], st["sigF"], st["sigH"], st["sigT"],
                         EMsteps=steps, atol=-np.inf)
        errs = dict(loglik=relmax(xa["loglik"], o["loglik"]), W=relmax(xa["W"], o["W"]), C=relmax(xa["C"], o["C"]),
                    B=relmax(xa["B"], o["B"]), sigT=relmax(xa["sigT"], o["sigT"]), sigE=abs(xa["sigE"] / o["sigE"] - 1),
                    vs_nccl=relmax(xa["W"], na["W"]), repeat=relmax(xa["W"], xb["W"]))
        from oracle import ppls_ref as R, ppls_seq
        so = ppls_seq.ppls(X, Y, 3, 15, 1e-4, "equal")
        errs["seq_steps"] = 0.0 if np.array_equal(seq["Other_output"]["Number_steps"], so["Number_steps"]) else 1.0
        errs["seq_W"] = relmax(seq["W"], so["W"]); errs["seq_C"] = relmax(seq["C"], so["C"])
        errs["seq_B"] = relmax(seq["B"], so["B"]); errs["seq_sig"] = float(np.max(np.abs(seq["sig"] / so["sig"] - 1)))
        errs["seq_loglik"] = max(relmax(t, so["logvalues"][i][:t.size]) for i, t in enumerate(seq["Other_output"]["logvalues"]))
        # PPLS_simult with its default start: the 20-step, 2-component sequential fit, then the loop
        s2 = ppls_seq.ppls(X, Y, 2, 20, 1e-4, "equal")
        W0, C0, B0, T0 = R.normalise(s2["W"], s2["C"], s2["B"], s2["sig"][:, 3])
        o2 = R.ppls_simult_loop(X, Y, W0, C0, B0, s2["sig"][1, 0], s2["sig"][1, 1], s2["sig"][1, 2], T0, EMsteps=6, atol=-np.inf)
        sim = seq["simult_default_start"]
        errs["simult_default_W"] = relmax(sim["estimates"]["W"], o2["estimates"]["W"])
        errs["simult_default_loglik"] = relmax(sim["loglik"], o2["loglik"])
        for gi, bres in enumerate(seq["batched"]):       # start 0 of the batched runs is the single fit above
            errs[f"batched{gi}_W"] = relmax(bres[0]["loop_state"]["W"], xa["W"])
            errs[f"batched{gi}_loglik"] = relmax(bres[0]["loglik"], xa["loglik"])
        from oracle import ppls_meta
        cuts = seq["meta_cuts"]
        mo = ppls_meta.meta_pplsi(X, Y, [int(c) for c in np.diff(cuts)], 8, -np.inf)
        errs["meta_W"] = relmax(seq["meta"]["W"], mo["W"]); errs["meta_C"] = relmax(seq["meta"]["C"], mo["C"])
        errs["meta_params"] = max(abs(seq["meta"]["params"][j][k] / mo["params"][j][k] - 1) for j in range(3)
                                  for k in ("B_T", "sigX", "sigY", "sigH", "sigT"))
        errs["meta_log"] = relmax(seq["meta"]["log"], np.asarray(mo["log"])[:seq["meta"]["log"].shape[0]])
        ok = (same_ranks and errs["loglik"] < 1e-10 and errs["meta_log"] < 1e-10 and errs["seq_loglik"] < 1e-10 and errs["simult_default_loglik"] < 1e-10
              and all(v < 1e-8 for v in errs.values()) and errs["repeat"] == 0.0 and errs["vs_nccl"] < 1e-12)
        # (the NCCL form reduces the CTAs' partial vectors in another fixed order than the fused tail does, so the two agree to
        # rounding, not bit for bit; every form is bit-reproducible run to run and bit-identical across ranks)
        if os.environ.get("PPLS_DUMP"):      # for the single-process multi-device test: must agree bit for bit
            np.savez(os.environ["PPLS_DUMP"], W=xa["W"], C=xa["C"], B=xa["B"], sigT=xa["sigT"], loglik=xa["loglik"],
                     sig=np.array([xa["sigE"], xa["sigF"], xa["sigH"]]))
        print("ROWSHARD", "OK" if ok else "FAIL", world, "ranks", (N, p, q, r), "same_on_all_ranks", same_ranks,
              {k: float(f"{v:.2e}") for k, v in errs.items()}, flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.stdout.flush()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
