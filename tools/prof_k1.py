"""Short run for ncu: a few passes of the streaming kernel on a resident synthetic shard.
Usage: python tools/prof_k1.py [workload] [rows] [passes]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from ppls_b200 import Context, Data, Fit, synth
from bench import WORKLOADS

wl = sys.argv[1] if len(sys.argv) > 1 else "c3"
N, p, q, r = WORKLOADS[wl]
if len(sys.argv) > 2 and int(sys.argv[2]) > 0:
    N = int(sys.argv[2])
passes = int(sys.argv[3]) if len(sys.argv) > 3 else 4
ctx = Context(0)
truth = synth.truth_params(p, q, r); start = synth.start_params(truth, 1)
data = Data.synthetic(ctx, N, p, q, truth, seed=1)
f = Fit(data, start, r, EMsteps=passes + 4, atol=-np.inf, timing=True, check_every=100)
f.iterate(passes)
t = f.timing()
print("k1 ms/launch", t["stream_ms"] / t["stream_launches"], "GB/s", 8 * N * (p + q) / (t["stream_ms"] / t["stream_launches"]) / 1e6)
f.close(); data.close(); ctx.close()
