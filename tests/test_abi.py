"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports exactly the
symbols include/ppls_b200.h declares, and refuses to compute without a GPU (no fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest

from ppls_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "ppls_b200.h")


def _declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ppls_[a-z_0-9]+)\s*\(", src)))


def test_header_and_binding_agree():
    decl = _declared_functions()
    assert len(decl) >= 19
    assert sorted(_lib.SIGNATURES) == decl


def test_library_exports_every_declared_symbol():
    L = ctypes.CDLL(_lib.LIB_PATH)
    for name in _declared_functions():
        assert hasattr(L, name), name
    assert b"sm_100a" in _lib.lib().ppls_version()


def test_struct_sizes_match_header():
    # ppls_opts: 4 x int32, double, 2 x uint32; ppls_params: 4 pointers + 3 doubles; ppls_timing: 10 x 8 bytes
    assert ctypes.sizeof(_lib.Opts) == 32
    assert ctypes.sizeof(_lib.Params) == 4 * ctypes.sizeof(ctypes.c_void_p) + 24
    assert ctypes.sizeof(_lib.Timing) == 80


def test_header_cites_reference_lines():
    src = open(HEADER).read()
    for cite in ("EM_W_multi.R:758-807", "loglC.cpp:319-338", "RcppExports.cpp:79-99", "RcppExports.R:32-34"):
        assert cite in src


def test_no_cpu_fallback():
    """Without a CUDA device every compute entry point fails loudly with PPLS_E_CUDA."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu tests")
    h = ctypes.c_void_p()
    rc = _lib.lib().ppls_ctx_create(ctypes.byref(h), 0)
    assert rc == _lib.PPLS_E_CUDA
    assert b"no CPU fallback" in _lib.lib().ppls_last_error()
    from ppls_b200 import PPLS_simult, PPLSError
    with pytest.raises(PPLSError):
        PPLS_simult(np.zeros((8, 3)), np.zeros((8, 2)), 1)


def test_product_never_imports_oracle():
    """oracle/ is test infrastructure: nothing under ppls_b200/ may import, load or link it."""
    pkg = os.path.join(ROOT, "ppls_b200")
    pat = re.compile(r"from\s+oracle|import\s+oracle|libppls_oracle|oracle[./]ppls_|cbind")
    for dirpath, _, files in os.walk(pkg):
        if os.path.basename(dirpath) in ("build", "__pycache__"):
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                assert not pat.search(open(os.path.join(dirpath, f)).read()), (dirpath, f)
