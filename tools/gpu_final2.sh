#!/bin/bash
# the driver's N = 1 commands with the final code (reference arm first, as the driver runs them), then the launch list
set -u
OUT=gpurun_out/final2
mkdir -p $OUT
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_reference.json 2> $OUT/bench_reference.err; echo "reference rc=$?"
timeout 1500 python bench.py > $OUT/bench_default.json 2> $OUT/bench_default.err; echo "bench rc=$?"; wc -l $OUT/bench_default.json
python - <<PY
import json
d = json.loads(open("$OUT/bench_default.json").read().strip().splitlines()[-1])
print("C3", round(d["value"], 2), "it/s frac", round(d["roofline"]["frac"], 3), "of nominal", round(d["roofline"]["frac_of_nominal_8000"], 3), "e2e", round(d["e2e"]["value"], 2), "launches", d["gpu_launches"], d["clocks"])
for k, v in d.get("other_workloads", {}).items():
    print(k, {kk: (round(vv, 4) if isinstance(vv, float) else vv) for kk, vv in v.items() if kk in ("it_per_s", "ms_per_step", "frac", "tail_ms_per_step", "error")})
PY
CMD="python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --rows 2000000"
$CMD > $OUT/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"em_stream|wide|em_update|bprojmu|baccum|bfinal|em_update_batch" -c 120 --csv --log-file $OUT/launches_default.csv $CMD > $OUT/ncu_launch.log 2>&1
echo "launch list rc=$?"
