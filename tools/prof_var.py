"""Time the variances path (Gram kernel, X'mu pass, inverse) on synthetic resident data.
Usage: python tools/prof_var.py [n p q r]   (prints the library's own event timings, PPLS_VAR_VERBOSE)"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["PPLS_VAR_VERBOSE"] = "1"
from ppls_b200 import Context, Data, Fit, synth, variances   # noqa: E402

n, p, q, r = (int(a) for a in sys.argv[1:5]) if len(sys.argv) >= 5 else (1_000_000, 1000, 200, 3)
ctx = Context(0)
truth = synth.truth_params(p, q, r)
d = Data.synthetic(ctx, n, p, q, truth, seed=7)
f = Fit(d, synth.start_params(truth, 1), r, 3, -np.inf, "SVD")
f.run()
res = f.result(want_mu=True)
f.close()
fit = dict(Expectations=dict(mu_T=res["mu_T"], mu_U=res["mu_U"], Ctt=np.diag(res["Ctt"]) if np.ndim(res["Ctt"]) == 1 else res["Ctt"],
                             Cuu=np.diag(res["Cuu"]) if np.ndim(res["Cuu"]) == 1 else res["Cuu"]),
           estimates=dict(sigE=res["sigE"]))
for rep in range(2):
    t0 = time.perf_counter()
    out = variances(fit, None, "X", dev_data=d, matrices=False)
    print("variances(X) wall %.3f s, seLoad[0] =" % (time.perf_counter() - t0), out["seLoad"][0], flush=True)
d.close(); ctx.close()
