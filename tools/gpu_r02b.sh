#!/bin/bash
# Round 2, session B (1 GPU): fused tail (one launch per EM iteration) against the three-kernel form.
set -u
OUT=gpurun_out/r02b
mkdir -p $OUT
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log
tail -15 $OUT/pytest_gpu.log
for F in 1 0; do
  for WL in c2 c3; do
    PPLS_FUSE_TAIL=$F timeout 600 python bench.py --workload $WL --steps 30 --warmup 5 --no-cpu --no-e2e --no-others > $OUT/bench_${WL}_fuse$F.json 2> $OUT/bench_${WL}_fuse$F.err
    echo "bench $WL fuse=$F rc=$?"
    python - <<PY
import json
try:
    d = json.load(open("$OUT/bench_${WL}_fuse$F.json"))
    print("$WL fuse=$F", round(d["value"], 2), "it/s", round(d["ms_per_step"], 4), "ms/step kernel", round(d["roofline"]["kernel_ms"], 4), "tail", round(d["config"]["tail_ms_per_step"], 4), "launches", d["gpu_launches"], "frac", round(d["roofline"]["frac"], 3), d["clocks"]["sm_mhz"])
except Exception as e:
    print("no line:", e)
PY
    tail -3 $OUT/bench_${WL}_fuse$F.err
  done
done
# KEEP form of the streaming kernel (tile read from shared memory once)
PPLS_K1_KEEP=1 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -m gpu -x -q > $OUT/pytest_keep.log 2>&1; echo "pytest KEEP rc=$?"; tail -3 $OUT/pytest_keep.log
for WL in c3 c2; do
  PPLS_K1_KEEP=1 PPLS_K1_VERBOSE=1 timeout 600 python bench.py --workload $WL --steps 30 --warmup 5 --no-cpu --no-e2e --no-others > $OUT/bench_${WL}_keep.json 2> $OUT/bench_${WL}_keep.err
  echo "bench $WL keep rc=$?"
  python - <<PY
import json
try:
    d = json.load(open("$OUT/bench_${WL}_keep.json"))
    print("$WL keep", round(d["value"], 2), "it/s", round(d["ms_per_step"], 4), "ms/step kernel", round(d["roofline"]["kernel_ms"], 4), "tail", round(d["config"]["tail_ms_per_step"], 4), "frac", round(d["roofline"]["frac"], 3), d["clocks"]["sm_mhz"])
except Exception as e:
    print("no line:", e)
PY
  grep "K1 plan" $OUT/bench_${WL}_keep.err | head -2
done
