"""Timing experiment driver for the batched path: a few passes of C5 whose numerical outcome is ignored
(PPLS_BGEMM_DBG switches the copies or the DMMAs off, so the fit is garbage); always exits 0 so that ncu runs."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ppls_b200 import Context, Data, PPLS_simult_batched, synth
torch.cuda.set_device(0)
ctx = Context(0)
N, p, q, r = 100_000, 200, 50, 2
truth = synth.truth_params(p, q, r)
data = Data.synthetic(ctx, N, p, q, truth, seed=5)
inits = [synth.start_params(truth, seed=1000 + s) for s in range(64)]
try:
    PPLS_simult_batched(None, None, r, inits, EMsteps=6, atol=-float("inf"), data=data)
    print("ok")
except Exception as e:
    print("ignored:", str(e)[:100])
os._exit(0)
