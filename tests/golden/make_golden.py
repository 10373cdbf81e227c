"""Generate the golden vectors of tests/golden/*.npz from the CPU oracle.

Run in the build container:  python tests/golden/make_golden.py
The reference has no tests or recorded outputs for PPLS_simult and cannot be executed here
(no R), so these vectors pin OUR restatement (oracle/ppls_ref.py), which tests/test_oracle.py
defends with the dense cross-checks.  Inputs come from the deterministic generator of
oracle/ppls_oracle.c and are stored in the fixture so that the GPU tests need nothing else.
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import cbind, ppls_ref as R  # noqa: E402
from ppls_b200 import synth  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

CASES = [
    # name, (N, p, q, r), (sigE, sigF, sigH), EMsteps, atol, type, seed
    ("c1_sig05_1000", (500, 20, 10, 2), (0.05, 0.05, 0.05), 1000, -np.inf, "SVD", 20261104),
    ("c1_wellcond_1000", (500, 20, 10, 2), (0.5, 0.5, 0.3), 1000, -np.inf, "SVD", 20261104),
    ("c1_break_rule", (500, 20, 10, 2), (0.5, 0.5, 0.3), 400, 1e-2, "SVD", 20261105),
    ("r1_p_lt_q", (203, 7, 13, 1), (0.5, 0.5, 0.3), 25, -np.inf, "SVD", 11),
    ("r5_odd_shapes", (333, 31, 17, 5), (0.3, 0.2, 0.1), 40, -np.inf, "SVD", 12),
    ("r3_qr", (257, 40, 24, 3), (0.4, 0.3, 0.2), 30, -np.inf, "QR", 13),
    ("r8_max", (129, 36, 20, 8), (0.3, 0.3, 0.2), 15, -np.inf, "SVD", 14),
    ("r3_wide_row", (97, 300, 90, 3), (0.5, 0.4, 0.3), 20, -np.inf, "SVD", 15),
    ("one_step", (64, 9, 5, 2), (0.5, 0.5, 0.3), 1, 1e-4, "SVD", 16),
]
CHECKPOINTS = (1, 2, 5, 10, 100, 1000)


def main():
    L = cbind.load()
    for name, (N, p, q, r), sig, steps, atol, typ, seed in CASES:
        truth = synth.truth_params(p, q, r, *sig)
        X, Y = cbind.gen_data(L, seed, 0, N, truth)
        st = synth.start_params(truth, seed=seed % 1000)
        # the reference normalises its start values before the loop (EM_W_multi.R:773-778)
        W0, C0, B0, T0 = R.normalise(st["W"], st["C"], st["B"], st["sigT"])
        ck = {}

        def rec(i, s):
            if i in CHECKPOINTS:
                ck[f"ck{i}_W"] = s["W"].copy(); ck[f"ck{i}_C"] = s["C"].copy()
                ck[f"ck{i}_B"] = np.diag(s["B"]).copy(); ck[f"ck{i}_sigT"] = np.diag(s["sigT"]).copy()
                ck[f"ck{i}_sig"] = np.array([s["sigE"], s["sigF"], s["sigH"]])

        out = R.ppls_simult_loop(X, Y, W0, C0, B0, st["sigE"], st["sigF"], st["sigH"], T0, EMsteps=steps, atol=atol,
                                 type=typ, record=rec)
        E, est, ls = out["Expectations"], out["estimates"], out["loop_state"]
        np.savez_compressed(
            os.path.join(HERE, name + ".npz"),
            X=X, Y=Y, seed=seed, r=r, EMsteps=steps, atol=atol, type=typ,
            truth_W=truth["W"], truth_C=truth["C"], truth_B=truth["B"], truth_sigT=truth["sigT"], truth_sig=np.array(sig),
            init_W=W0, init_C=C0, init_B=np.diag(B0), init_sigT=np.diag(T0),
            init_sig=np.array([st["sigE"], st["sigF"], st["sigH"]]),
            loglik=out["loglik"],
            est_W=est["W"], est_C=est["C"], est_B=np.diag(est["B"]), est_sigT=np.diag(est["sigT"]),
            est_sig=np.array([est["sigE"], est["sigF"], est["sigH"]]),
            loop_W=ls["W"], loop_C=ls["C"], loop_B=np.diag(ls["B"]), loop_sigT=np.diag(ls["sigT"]),
            mu_T=E["mu_T"], mu_U=E["mu_U"], Ctt=np.diag(E["Ctt"]), Cuu=np.diag(E["Cuu"]), Cut=np.diag(E["Cut"]),
            Cee=E["Cee"][0, 0], Cff=E["Cff"][0, 0], Chh=E["Chh"], **ck)
        print(f"{name}: steps={len(out['loglik'])} loglik[-1]={out['loglik'][-1]:.6f}")


if __name__ == "__main__":
    main()
