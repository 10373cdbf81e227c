#!/bin/bash
# GPU test suite only (usage under gpurun: bash tools/gpu_tests.sh <tag>)
set -u
OUT=gpurun_out/${1:-tests}
mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log
tail -40 $OUT/pytest_gpu.log | cut -c1-700
