#!/bin/bash
# session M (1 GPU): the round's evidence: default bench line, reference arm, launch list, full ncu captures
set -u
OUT=gpurun_out/r02m
mkdir -p $OUT
timeout 1500 python bench.py > $OUT/bench_default.json 2> $OUT/bench_default.err; echo "bench rc=$?"
python - <<PY
import json
d = json.load(open("$OUT/bench_default.json"))
print("C3", round(d["value"], 2), "it/s frac", round(d["roofline"]["frac"], 3), "of nominal", round(d["roofline"]["frac_of_nominal_8000"], 3), "e2e", d["e2e"]["value"], "launches", d["gpu_launches"], d["clocks"])
for k, v in d.get("other_workloads", {}).items():
    print(k, {kk: (round(vv, 4) if isinstance(vv, float) else vv) for kk, vv in v.items() if kk in ("value", "ms_per_step", "frac", "tail_ms_per_step", "achieved", "unit")})
print("cpu", d["cpu_baseline"])
PY
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_reference.json 2> $OUT/bench_reference.err; echo "reference rc=$?"; cut -c1-400 $OUT/bench_reference.json
CMD="python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --rows 2000000"
$CMD > $OUT/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_default.csv $CMD > $OUT/ncu_launch.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:em_stream_kernel -s 4 -c 1 -f -o $OUT/k1_full python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --no-others --rows 2000000 > $OUT/ncu_k1.log 2>&1
echo "k1 full rc=$?"
