#!/bin/bash
# session O (8 GPUs): C2 with the fused tail's clock breakdown, then the driver's default command at N = 8
set -u
N=${1:-8}
OUT=gpurun_out/r02o
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
PPLS_TAIL_VERBOSE=1 timeout 600 $TR --master-port 29571 bench.py --gpus $N --workload c2 --steps 100 --warmup 10 --no-cpu --no-e2e --no-others --no-parity > $OUT/bench_c2_${N}gpu.json 2> $OUT/bench_c2_${N}gpu.err; echo "c2 rc=$?"
grep "fused tail" $OUT/bench_c2_${N}gpu.err | tail -4
PPLS_FUSE_TAIL=0 timeout 600 $TR --master-port 29572 bench.py --gpus $N --workload c2 --steps 100 --warmup 10 --no-cpu --no-e2e --no-others --no-parity > $OUT/bench_c2_${N}gpu_fuse0.json 2> $OUT/bench_c2_${N}gpu_fuse0.err; echo "c2 fuse0 rc=$?"
timeout 2400 $TR --master-port 29573 bench.py --gpus $N --steps 20 --warmup 3 > $OUT/bench_${N}gpu.json 2> $OUT/bench_${N}gpu.err; echo "bench rc=$?"
python - <<PY
import json
for f in ("bench_c2_${N}gpu", "bench_c2_${N}gpu_fuse0", "bench_${N}gpu"):
    try:
        txt = open("$OUT/" + f + ".json").read()
        d = json.loads([l for l in txt.splitlines() if l.startswith('{"metric')][-1])
    except Exception as e:
        print(f, "no line", e); continue
    print(f, round(d["value"], 2), "it/s ms/step", round(d["ms_per_step"], 4), "tail", d["config"].get("tail_ms_per_step"), "frac", round(d["roofline"]["frac"], 3))
    print("   per rank [step, k1, tail]:", d["config"].get("per_rank_ms_step_k1_tail"))
    for k, v in d.get("other_workloads", {}).items():
        print("   ", k, {kk: (round(vv, 4) if isinstance(vv, float) else vv) for kk, vv in v.items() if kk in ("value", "ms_per_step", "frac", "tail_ms_per_step")})
    if "parity" in d: print("    parity", d["parity"].get("pass"), d["parity"].get("bit_identical_ranks"))
    if "e2e" in d and d["e2e"]: print("    e2e", d["e2e"].get("value"), "single process", (d.get("e2e_single_process") or {}).get("value"))
PY
nvidia-smi --query-gpu=index,name,clocks.sm,power.draw --format=csv > $OUT/gpus.csv
