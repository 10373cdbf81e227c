#!/bin/bash
# Round 2, session C (1 GPU): fused tail v2 (CTA 0 off the row path) + KEEP form of the streaming kernel.
set -u
OUT=gpurun_out/r02c
mkdir -p $OUT
line() {  # $1 = json file, $2 = label
python - <<PY
import json
try:
    d = json.load(open("$1"))
    print("$2", round(d["value"], 2), "it/s", round(d["ms_per_step"], 4), "ms/step kernel", round(d["roofline"]["kernel_ms"], 4), "tail", round(d["config"]["tail_ms_per_step"], 4), "launches", d["gpu_launches"], "frac", round(d["roofline"]["frac"], 3), d["clocks"]["sm_mhz"], "probe", d["roofline"].get("read_only_probe_gbs"))
except Exception as e:
    print("$2 no line:", e)
PY
}
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log
tail -4 $OUT/pytest_gpu.log
PPLS_K1_KEEP=1 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -m gpu -x -q > $OUT/pytest_keep.log 2>&1; echo "pytest KEEP rc=$?"; tail -3 $OUT/pytest_keep.log
for WL in c2 c3; do
  for K in 0 1; do
    PPLS_K1_KEEP=$K PPLS_K1_VERBOSE=1 timeout 600 python bench.py --workload $WL --steps 30 --warmup 5 --no-cpu --no-e2e --no-others > $OUT/bench_${WL}_keep$K.json 2> $OUT/bench_${WL}_keep$K.err
    echo "bench $WL keep=$K rc=$?"; line $OUT/bench_${WL}_keep$K.json "$WL keep=$K"
    grep "K1 plan" $OUT/bench_${WL}_keep$K.err | head -1
  done
done
# KEEP with other warp counts / tiles at C3
for CFG in "5 4 4" "5 4 5" "6 4 4" "8 4 4" "5 2 6"; do
  set -- $CFG
  PPLS_K1_KEEP=1 PPLS_K1_WARPS=$1 PPLS_K1_T=$2 PPLS_K1_STAGES=$3 PPLS_K1_VERBOSE=1 timeout 600 python bench.py --workload c3 --rows 4000000 --steps 30 --warmup 5 --no-cpu --no-e2e --no-others > $OUT/b.json 2> $OUT/b.err
  line $OUT/b.json "c3/4M keep warps=$1 T=$2 stages=$3"; grep "K1 plan" $OUT/b.err | head -1
done
PPLS_K1_KEEP=0 timeout 600 python bench.py --workload c3 --rows 4000000 --steps 30 --warmup 5 --no-cpu --no-e2e --no-others > $OUT/b.json 2> $OUT/b.err
line $OUT/b.json "c3/4M keep=0"
