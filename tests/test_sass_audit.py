"""Static check of the compiled kernels: no shared-memory load may be outstanding when a TMA-filled stage is released.

Background ("A hardware lesson"): ptxas orders an mbarrier arrive after the ISSUE of earlier shared-memory loads, not
after their return; in the single-pass wide kernel a refill of the stage overtook such loads and 15-90 % of runs had wrong
32-byte sectors.  The fix is a data dependency (zero_once_known, em_stream_kernel.cuh); tools/sass_arrive_audit.py reads the
SASS of an object file and reports every stage-release arrive with unconsumed LDS destinations.  This test keeps the kernels
that stream through shared-memory rings clean.  Needs the in-tree build (python -m ppls_b200.build) and cuobjdump; no GPU."""
import os
import shutil
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJECTS = ["em_wide3.o", "em_wide2.o", "em_wide.o", "em_stream_inst_r3.o"]


def test_no_load_outstanding_at_a_stage_release():
    if shutil.which("cuobjdump") is None:
        pytest.skip("cuobjdump not on PATH")
    objs = [os.path.join(ROOT, "ppls_b200", "build", o) for o in OBJECTS]
    if not all(os.path.exists(o) for o in objs):
        pytest.skip("object files not built in-tree (python -m ppls_b200.build)")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "sass_arrive_audit.py"), *objs],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == len(objs) and all("no outstanding loads at any arrive" in l for l in lines), r.stdout[-3000:]
