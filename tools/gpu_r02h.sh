#!/bin/bash
# Round 2, session H (8 GPUs): the driver's multi-GPU bench line (C3 + C2/C4/C5 + parity + single-process e2e), C2 fused vs three-kernel form.
set -u
OUT=gpurun_out/r02h
mkdir -p $OUT
nvidia-smi --query-gpu=index,name --format=csv > $OUT/gpus.csv
N=${1:-8}
SECONDS=0
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29610 bench.py --gpus $N --steps 20 --warmup 3 > $OUT/bench_${N}gpu.json 2> $OUT/bench_${N}gpu.err
echo "bench ${N}gpu rc=$? in ${SECONDS}s"; tail -c 4500 $OUT/bench_${N}gpu.json; tail -4 $OUT/bench_${N}gpu.err
for F in 1 0; do
SECONDS=0
PPLS_FUSE_TAIL=$F timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --workload c2 --steps 100 --warmup 10 --no-e2e --no-others --no-parity > $OUT/bench_c2_${N}gpu_fuse$F.json 2> $OUT/bench_c2_${N}gpu_fuse$F.err
echo "bench c2 ${N}gpu fuse=$F rc=$? in ${SECONDS}s"
python - <<PY
import json
for line in open("$OUT/bench_c2_${N}gpu_fuse$F.json"):
    if line.startswith("{"):
        d = json.loads(line)
        print("c2 fuse=$F", round(d["value"], 1), "it/s", round(d["ms_per_step"], 4), "ms/step kernel", round(d["roofline"]["kernel_ms"], 4), "tail", round(d["config"]["tail_ms_per_step"], 4), d["config"]["per_rank_ms_step_k1_tail"][:3])
PY
done
