#!/bin/bash
# session K (1 GPU): single-pass wide path (em_wide3.cu): parity, C4 bench, launch list
set -u
OUT=gpurun_out/r02k
mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_seq.py -m gpu -x -q -k "wide" > $OUT/pytest_wide.log 2>&1; echo "pytest wide rc=$?" | tee -a $OUT/pytest_wide.log
tail -12 $OUT/pytest_wide.log | cut -c1-1200
for CFG in "1 3 5" "1 4 5" "1 2 5" "1 4 4" "0 3 4"; do
  set -- $CFG
  PPLS_WIDE_SINGLE=$1 PPLS_WIDE_LAG=$2 PPLS_WIDE_STAGES=$3 PPLS_K1_VERBOSE=1 timeout 300 python bench.py --workload c4 --steps 100 --warmup 10 --no-cpu --no-e2e --no-others > $OUT/bench_c4_$1_$2_$3.json 2> $OUT/bench_c4_$1_$2_$3.err
  echo "bench c4 single=$1 lag=$2 stages=$3 rc=$?"
  python - <<PY
import json
try:
    d = json.load(open("$OUT/bench_c4_$1_$2_$3.json"))
    print("  ", round(d["value"], 1), "it/s", round(d["ms_per_step"], 4), "ms/step kernels", round(d["roofline"]["kernel_ms"], 4), "tail", round(d["config"]["tail_ms_per_step"], 4), "frac", round(d["roofline"]["frac"], 3), d["clocks"]["sm_mhz"], "launches", d["gpu_launches"])
except Exception as e:
    print("no line:", e)
PY
  grep "wide plan" $OUT/bench_c4_$1_$2_$3.err | tail -1
done
PPLS_WIDE_HINT=0 timeout 300 python bench.py --workload c4 --steps 100 --warmup 10 --no-cpu --no-e2e --no-others 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('no L2 hints:', round(d['value'],1), 'it/s kernels', round(d['roofline']['kernel_ms'],4))"
CMD="python bench.py --workload c4 --steps 6 --warmup 3 --no-cpu --no-e2e --no-others"
$CMD > $OUT/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum --clock-control none -k regex:wide3 -c 6 --csv --log-file $OUT/c4_wide3_metrics.csv $CMD > $OUT/ncu_launch.log 2>&1
echo "ncu rc=$?"
tail -8 $OUT/c4_wide3_metrics.csv | cut -c1-400
PPLS_TAIL_VERBOSE=1 timeout 300 python bench.py --workload c2 --steps 50 --warmup 5 --no-cpu --no-e2e --no-others 2>&1 >/dev/null | grep "fused tail" | tail -2
