"""Sweep the streaming-kernel launch parameters (M, T, stages, CTAs/SM) on one resident dataset.
Usage: python tools/tune_k1.py c3 [rows]   (GPU box)"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from ppls_b200 import Context, Data, Fit, synth, PPLSError
from bench import WORKLOADS

wl = sys.argv[1] if len(sys.argv) > 1 else "c3"
N, p, q, r = WORKLOADS[wl]
if len(sys.argv) > 2:
    N = int(sys.argv[2])
P = int(sys.argv[3]) if len(sys.argv) > 3 else 6
ctx = Context(0)
truth = synth.truth_params(p, q, r); start = synth.start_params(truth, 1)
t0 = time.time(); data = Data.synthetic(ctx, N, p, q, truth, seed=1); print("generate", round(time.time() - t0, 2), "s", flush=True)
grid = []
for Wn in (4, 8):
    for T in (4, 8):
        for st in (4, 5):
            grid.append((Wn, T, st))
res = []
for Wn, T, st in grid:
    os.environ.update(PPLS_K1_WARPS=str(Wn), PPLS_K1_T=str(T), PPLS_K1_STAGES=str(st))
    try:
        f = Fit(data, start, r, EMsteps=P + 10, atol=-np.inf, timing=True, check_every=100)
    except PPLSError as e:
        continue
    f.iterate(2); a = f.timing(); f.iterate(P); b = f.timing(); f.close()
    ms = (b["stream_ms"] - a["stream_ms"]) / (b["stream_launches"] - a["stream_launches"])
    loop = (b["loop_ms"] - a["loop_ms"]) / P
    gbs = 8 * N * (p + q) / ms / 1e6
    res.append((gbs, Wn, T, st, ms, loop))
    print(f"warps={Wn} T={T} stages={st}: k1 {ms:.3f} ms  loop/iter {loop:.3f} ms  {gbs:.0f} GB/s", flush=True)
res.sort(reverse=True)
print("best:", res[:5])
