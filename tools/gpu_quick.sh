#!/bin/bash
# Quick GPU check: parity tests + K1 timing at C3/C2 widths.  Usage: bash tools/gpu_quick.sh tag
OUT=gpurun_out/${1:-quick}
mkdir -p $OUT
python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 $OUT/pytest_gpu.log
export PPLS_K1_VERBOSE=1
python tools/prof_k1.py c3 3000000 6 2>&1 | tail -3
python tools/prof_k1.py c2 1000000 6 2>&1 | tail -3
