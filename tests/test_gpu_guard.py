"""Out-of-bounds writes: every kernel family under the engine's guarded allocator (PPLS_GUARD=1, capi.cu).

compute-sanitizer is closed on the pool these tests run on (profiles/r02/sanitizer/memcheck_refused.log), so the engine
checks itself: with PPLS_GUARD=1 each device allocation has 4 KB of pattern bytes on either side, verified when it is freed;
a changed byte aborts the process.  tools/san_cases.py drives the fused one-launch path, the update kernels, the three wide-row
forms, the batched paths, sequential / meta fits, variances, orth and the allocation-free orth on small ragged cases."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _run(extra_env, *args):
    env = dict(os.environ, PPLS_GUARD="1", **extra_env)
    return subprocess.run([sys.executable, os.path.join(ROOT, "tools", "san_cases.py"), *args], env=env, cwd=ROOT,
                          capture_output=True, text=True, timeout=900)


def test_no_write_outside_any_device_allocation():
    r = _run({})
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "orth noalloc" in r.stdout and "PPLS_GUARD" not in r.stderr


def test_no_write_outside_any_allocation_on_two_devices():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    r = _run({}, "--multi")
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "multi QR" in r.stdout and "PPLS_GUARD" not in r.stderr


def test_the_guard_detects_a_damaged_byte():
    r = _run({"PPLS_GUARD_SELFTEST": "1"})
    assert r.returncode != 0
    assert "PPLS_GUARD: write 3 bytes PAST the end" in r.stderr
