#!/bin/bash
# Round 2, session G (1 GPU): wide-row path on the FP64 tensor pipe (em_wide2.cu) against the DFMA form.
set -u
OUT=gpurun_out/r02g
mkdir -p $OUT
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log
tail -15 $OUT/pytest_gpu.log | cut -c1-600
for D in 1 0; do
  PPLS_WIDE_DMMA=$D PPLS_K1_VERBOSE=1 timeout 600 python bench.py --workload c4 --steps 50 --warmup 5 --no-cpu --no-e2e --no-others > $OUT/bench_c4_dmma$D.json 2> $OUT/bench_c4_dmma$D.err
  echo "bench c4 dmma=$D rc=$?"
  python - <<PY
import json
try:
    d = json.load(open("$OUT/bench_c4_dmma$D.json"))
    print("c4 dmma=$D", round(d["value"], 1), "it/s", round(d["ms_per_step"], 4), "ms/step passes", round(d["roofline"]["kernel_ms"], 4), "tail", round(d["config"]["tail_ms_per_step"], 4), "frac", round(d["roofline"]["frac"], 3), d["clocks"]["sm_mhz"], "launches", d["gpu_launches"])
except Exception as e:
    print("no line:", e)
PY
  grep "wide plan" $OUT/bench_c4_dmma$D.err | head -2
done
CMD="python bench.py --workload c4 --steps 6 --warmup 3 --no-cpu --no-e2e --no-others"
$CMD > $OUT/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/c4_launches.csv $CMD > $OUT/ncu_launch.log 2>&1
echo "ncu launches rc=$?"
python tools/launch_times.py $OUT/c4_launches.csv 2>/dev/null | tail -15
