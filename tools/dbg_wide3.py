"""Debug: C4 through the single-pass and the two-pass wide forms; where do they differ?"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ppls_b200 import Context, Data, Fit, synth

N, p, q, r = 2000, 50000, 1000, 5
ctx = Context(0)
truth = synth.truth_params(p, q, r, 0.5, 0.5, 0.3)
d = Data.synthetic(ctx, N, p, q, truth, seed=777)
st = synth.start_params(truth, 3)

def run(single, steps=1):
    os.environ["PPLS_WIDE_SINGLE"] = single
    f = Fit(d, st, r, steps, -np.inf, "SVD")
    f.run()
    res = f.result()
    f.close()
    return res

a = run("0")
nbad = 0
REPS = int(os.environ.get("REPS", "60"))
for rep in range(REPS):
    b = run("1")
    bad = False
    for k in ("W", "C"):
        x, y = np.asarray(a[k]), np.asarray(b[k])
        dlt = np.abs(x - y)
        if dlt.max() > 1e-9 * np.abs(x).max():
            bad = True
            percol = dlt.max(axis=1)
            npan = (len(percol) + 127) // 128
            pan = np.array([percol[i * 128:(i + 1) * 128].max() for i in range(npan)])
            top = np.argsort(-pan)[:3]
            t0 = int(top[0]); seg = percol[t0 * 128:(t0 + 1) * 128]
            print("rep", rep, k, "max diff", dlt.max(), "worst panels", [(int(t), float("%.2g" % pan[t])) for t in top],
                  "cols", np.where(seg > 10 * np.median(pan))[0][:24], "pos", t0 % 148, "slot", t0 // 148)
    if not bad and abs(a["loglik"][0] - b["loglik"][0]) > 1e-10 * abs(a["loglik"][0]):
        bad = True
        dm = np.abs(np.asarray(a["mu_T"]) - np.asarray(b["mu_T"])).max(axis=1)
        print("rep", rep, "loglik only", a["loglik"][0], b["loglik"][0], "rows off", np.where(dm > 1e-9)[0][:40])
    nbad += bad
print("bad", nbad, "of", REPS)
