"""ppls_b200 -- B200-native engine for the PPLS_simult EM loop (selbouhaddani/PPLS).

Only what the hot path needs: csrc/ (sm_100a CUDA kernels + the C ABI of
include/ppls_b200.h) and a host-side mirror of the reference's R interface (api.py).
"""
from .api import (Context, Data, Fit, PPLS, PPLS_simult, PPLS_simult_batched, default_context, default_init,  # noqa: F401
                  fconstraint, logl_W, meta_PPLSi, normalise, orth, simulC, variances)
from ._lib import PPLSError, lib  # noqa: F401

__all__ = ["Context", "Data", "Fit", "PPLS", "fconstraint", "PPLS_simult", "PPLS_simult_batched", "logl_W", "orth", "variances", "simulC", "meta_PPLSi", "default_context", "default_init", "normalise",
           "PPLSError", "lib"]
