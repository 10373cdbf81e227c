#!/bin/bash
# session N (N GPUs): the driver's multi-GPU commands: reference arm, then our arm, under torchrun
set -u
N=${1:-2}
OUT=gpurun_out/r02n
mkdir -p $OUT
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29561 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > $OUT/bench_reference_${N}gpu.json 2> $OUT/bench_reference_${N}gpu.err; echo "reference rc=$?"; cut -c1-300 $OUT/bench_reference_${N}gpu.json
timeout 2400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29562 bench.py --gpus $N --steps 20 --warmup 3 > $OUT/bench_${N}gpu.json 2> $OUT/bench_${N}gpu.err; echo "bench rc=$?"
python - <<PY
import json
d = json.load(open("$OUT/bench_${N}gpu.json"))
print("C3", round(d["value"], 2), "it/s frac", round(d["roofline"]["frac"], 3), "e2e", d["e2e"]["value"], "launches", d["gpu_launches"], d["clocks"])
for k, v in d.get("other_workloads", {}).items():
    print(k, {kk: (round(vv, 4) if isinstance(vv, float) else vv) for kk, vv in v.items() if kk in ("value", "ms_per_step", "frac", "tail_ms_per_step", "achieved", "sharding")})
print("parity", d.get("parity"))
print("e2e_single_process", d.get("e2e_single_process"))
PY
tail -3 $OUT/bench_${N}gpu.err | cut -c1-300
