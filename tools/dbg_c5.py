"""Timing experiments on the batched kernels (PPLS_BGEMM_DBG != 0 makes the results garbage and the fit fail: ignored here)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ppls_b200 import Context, Data, PPLS_simult_batched, synth

N, p, q, r, S = 100_000, 200, 50, 2, 64
ctx = Context(0)
truth = synth.truth_params(p, q, r)
d = Data.synthetic(ctx, N, p, q, truth, seed=5)
inits = [synth.start_params(truth, 10 + j) for j in range(S)]
try:
    PPLS_simult_batched(None, None, r, inits, EMsteps=6, atol=-np.inf, data=d)
except Exception as e:
    print("ignored:", str(e)[:80])
d.close(); ctx.close()
