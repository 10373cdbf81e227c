#!/bin/bash
# final check of the round (2 GPUs): the whole GPU suite, the no-alloc test alone, smoke(), then the driver's default bench commands at N = 1
set -u
OUT=gpurun_out/final
mkdir -p $OUT
bash tools/gpu_tests.sh final
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_reference.json 2> $OUT/bench_reference.err; echo "reference rc=$?"
timeout 1500 python bench.py > $OUT/bench_default.json 2> $OUT/bench_default.err; echo "bench rc=$?"
python - <<PY
import json
d = json.loads(open("$OUT/bench_default.json").read().strip().splitlines()[-1])
print("C3", round(d["value"], 2), "it/s frac", round(d["roofline"]["frac"], 3), "of nominal", round(d["roofline"]["frac_of_nominal_8000"], 3), "e2e", round(d["e2e"]["value"], 2), "launches", d["gpu_launches"], d["clocks"])
for k, v in d.get("other_workloads", {}).items():
    print(k, {kk: (round(vv, 4) if isinstance(vv, float) else vv) for kk, vv in v.items() if kk in ("it_per_s", "ms_per_step", "frac", "tail_ms_per_step", "error")})
print("cpu", d["cpu_baseline"]["value"], d["cpu_baseline"].get("fused_cpu_value"), d["cpu_baseline"]["cores"])
PY
wc -l $OUT/bench_default.json
