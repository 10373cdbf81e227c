"""Build libppls_b200.so in-tree with nvcc for sm_100a (no JIT cache: the .so travels with the repo)."""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libppls_b200.so")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-fno-strict-aliasing"]
# translation unit -> extra flags.  -fmad=false where bit-exact / reference-order arithmetic matters.
UNITS = {
    "em_stream.cu": ["-fmad=false"],
    # -fmad=false: the fused tail evaluates the closed-form coefficients (em_update_dev.cuh); the streaming loops
    # spell their FMAs out (fma()), so nothing hot depends on contraction
    **{f"em_stream_inst.cu@r{r}": [f"-DPPLS_INST_R={r}", "-fmad=false"] for r in range(1, 9)},
    "em_wide.cu": [],
    "em_wide2.cu": [],
    "em_wide3.cu": [],
    "em_update.cu": ["-fmad=false"],
    "em_seq.cu": ["-fmad=false"],
    "em_var.cu": [],
    "em_batch.cu": (["-DPPLS_HSTAGES=" + os.environ["PPLS_HSTAGES"]] if os.environ.get("PPLS_HSTAGES") else []),
    "data.cu": ["-fmad=false"],
    "probe.cu": [],
    "capi.cu": [],
}


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _stale(out: str, deps: list[str]) -> bool:
    if not os.path.exists(out):
        return True
    t = os.path.getmtime(out)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    nvcc = _nvcc()
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    headers = [os.path.join(CSRC, "common.cuh"), os.path.join(CSRC, "em_stream_kernel.cuh"),
               os.path.join(CSRC, "em_update_dev.cuh"), os.path.join(CSRC, "em_tail.cuh"),
               os.path.join(HERE, "..", "include", "ppls_b200.h")]
    objs = []
    procs = []
    for unit, extra in UNITS.items():
        src, _, tag = unit.partition("@")
        s = os.path.join(CSRC, src)
        o = os.path.join(objdir, src.replace(".cu", (("_" + tag) if tag else "") + ".o"))
        objs.append(o)
        if force or _stale(o, [s] + headers):
            cmd = [nvcc, *ARCH, *COMMON, *extra, "-c", s, "-o", o]
            if verbose:
                print(" ".join(cmd), file=sys.stderr)
            procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for cmd, pr in procs:
        out, _ = pr.communicate()
        if pr.returncode != 0:
            raise RuntimeError("nvcc failed: " + " ".join(cmd) + "\n" + out)
    if force or procs or _stale(LIB, objs):
        cmd = [nvcc, *ARCH, "-shared", "-o", LIB, *objs, "-ldl"]
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed: " + r.stdout)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
