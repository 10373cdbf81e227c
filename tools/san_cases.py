"""Small cases of every kernel family for compute-sanitizer (one --tool per gpurun call):
    compute-sanitizer --tool memcheck|racecheck|synccheck python tools/san_cases.py [--multi]
C1 golden (fused one-launch path, 40 steps), QR type (update kernel K2), the wide-row path on both storage layouts
(DFMA and DMMA two-pass forms, single-pass form), the batched paths (K1 packs and the two-GEMM form), sequential + meta fits, variances, orth,
the no-alloc orth; --multi: the row-sharded fit on a 2-device single-process context (peer exchange)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import init_of, load_golden
from ppls_b200 import (Context, Data, Fit, PPLS, PPLS_simult, PPLS_simult_batched, meta_PPLSi, orth, synth, variances)


def fit(ctx, X, Y, st, r, steps, typ="SVD"):
    d = Data.from_host(ctx, X, Y)
    f = Fit(d, st, r, steps, -np.inf, typ)
    f.run()
    res = f.result(want_mu=True)
    f.close(); d.close()
    return res


def main():
    multi = "--multi" in sys.argv
    g = load_golden("c1_wellcond_1000")
    st = init_of(g)
    if multi:
        ctx = Context.multi([0, 1])
        N, p, q, r = 6001, 301, 77, 3
        truth = synth.truth_params(p, q, r, 0.5, 0.5, 0.3)
        d = Data.synthetic(ctx, N, p, q, truth, seed=4242)
        for typ in ("SVD", "QR"):
            f = Fit(d, synth.start_params(truth, 3), r, 6, -np.inf, typ)
            f.run(); res = f.result(want_mu=True); f.close()
            print("multi", typ, res["loglik"][-1], flush=True)
        d.close(); ctx.close()
        return
    ctx = Context(0)
    res = fit(ctx, g["X"], g["Y"], st, 2, 40)
    print("c1 fused", abs(res["loglik"][-1] / g["loglik"][39] - 1), flush=True)
    # a shape with a full grid (fused tail across 148 CTAs) and ragged tiles
    N, p, q, r = 2501, 301, 77, 3
    truth = synth.truth_params(p, q, r, 0.5, 0.5, 0.3)
    d = Data.synthetic(ctx, N, p, q, truth, seed=7)
    for typ in ("SVD", "QR"):
        f = Fit(d, synth.start_params(truth, 3), r, 5, -np.inf, typ)
        f.run(); res = f.result(want_mu=True); f.close()
        print("grid", typ, res["loglik"][-1], flush=True)
    os.environ["PPLS_FUSE_TAIL"] = "0"
    f = Fit(d, synth.start_params(truth, 3), r, 4, -np.inf, "SVD"); f.run(); print("unfused", f.result(False)["loglik"][-1]); f.close()
    os.environ.pop("PPLS_FUSE_TAIL")
    fitres = PPLS_simult(None, None, r, EMsteps=3, atol=-np.inf, data=d, seed=1)
    print("default start", fitres["loglik"][-1], flush=True)
    Xh, Yh = d.export()
    v = variances(fitres, Xh, "X", ctx=ctx)
    print("variances", float(v["seLoad"].sum()), flush=True)
    inits = [synth.start_params(truth, 50 + j) for j in range(3)]
    for gemm in ("0", "1"):
        os.environ["PPLS_BATCH_GEMM"] = gemm
        b = PPLS_simult_batched(None, None, r, inits, EMsteps=3, atol=-np.inf, data=d)
        print("batched gemm=" + gemm, b[0]["loglik"][-1], flush=True)
    os.environ.pop("PPLS_BATCH_GEMM")
    s = PPLS(None, None, 2, 6, 1e-4, "random", data=d, seed=3, constraints=[dict(sigH=0.2), {}], loglikelihoods=True)
    print("sequential", s["B"], flush=True)
    m = meta_PPLSi(Xh, Yh, np.repeat([0, 1, 2], [1000, 800, 701]), 4, -np.inf, ctx=ctx)
    print("meta", m["Number_steps"], flush=True)
    d.close()
    w = load_golden("r3_wide_row")
    os.environ["PPLS_FORCE_WIDE"] = "1"
    for panel, dmma, single in (("0", "1", "0"), ("1", "0", "0"), ("1", "1", "0"), ("1", "1", "1")):
        os.environ["PPLS_PANEL"] = panel; os.environ["PPLS_WIDE_DMMA"] = dmma; os.environ["PPLS_WIDE_SINGLE"] = single
        res = fit(ctx, w["X"], w["Y"], init_of(w), 3, 6)
        print("wide panel=" + panel + " dmma=" + dmma + " single=" + single, abs(res["loglik"][-1] / w["loglik"][5] - 1), flush=True)
    for k in ("PPLS_FORCE_WIDE", "PPLS_PANEL", "PPLS_WIDE_DMMA", "PPLS_WIDE_SINGLE"):
        os.environ.pop(k)
    # the single-pass wide kernel with several panels per CTA and several row lanes
    for shape in ((300, 40000, 500, 5), (700, 6000, 500, 3)):
        N2, p2, q2, r2 = shape
        tr2 = synth.truth_params(p2, q2, r2, 0.5, 0.5, 0.3)
        d2 = Data.synthetic(ctx, N2, p2, q2, tr2, seed=9)
        f = Fit(d2, synth.start_params(tr2, 3), r2, 3, -np.inf, "SVD"); f.run()
        print("wide single-pass", shape, f.result(False)["loglik"][-1], flush=True); f.close(); d2.close()
    rng = np.random.default_rng(0)
    G = rng.standard_normal((1500, 4))
    for typ in ("SVD", "QR"):
        print("orth", typ, float(np.abs(orth(G, typ, ctx=ctx)).sum()), flush=True)
    import torch
    from ppls_b200 import _lib as L
    gt = torch.from_numpy(G).to("cuda:0"); ot = torch.empty_like(gt); torch.cuda.synchronize()
    L.check(L.lib().ppls_orth_noalloc(ctx._h, gt.data_ptr(), 1500, 4, 1, ot.data_ptr()))
    print("orth noalloc", float(ot.abs().sum()), flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
