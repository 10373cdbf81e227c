#!/bin/bash
# session I (N GPUs): multi-GPU tests + C2 with the per-CTA-flag exchange
set -u
N=${1:-2}
OUT=gpurun_out/r02i_$N
mkdir -p $OUT
if [ "$N" = "2" ]; then
timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "two_gpus or multi_device" > $OUT/pytest_mgpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_mgpu.log
tail -12 $OUT/pytest_mgpu.log | cut -c1-1500
fi
for WL in c2 c3; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --workload $WL --steps 100 --warmup 10 --no-e2e --no-others > $OUT/bench_${WL}.json 2> $OUT/bench_${WL}.err
echo "bench $WL rc=$?"
python - <<PY
import json
for line in open("$OUT/bench_${WL}.json"):
    if line.startswith("{"):
        d = json.loads(line)
        print("$WL", round(d["value"], 1), "it/s", round(d["ms_per_step"], 4), "ms/step kernel", round(d["roofline"]["kernel_ms"], 4), "tail", round(d["config"]["tail_ms_per_step"], 4), "scalar", round(d["config"]["tail_scalar_part_ms_per_step"], 4), d["config"]["per_rank_ms_step_k1_tail"][:2], d.get("parity", {}).get("pass"))
PY
tail -2 $OUT/bench_${WL}.err | cut -c1-300
done
timeout 300 python bench.py --workload c2 --steps 100 --warmup 10 --no-e2e --no-others --no-cpu > $OUT/bench_c2_1gpu.json 2>$OUT/bench_c2_1gpu.err
python - <<PY
import json
d = json.load(open("$OUT/bench_c2_1gpu.json"))
print("c2 1gpu", round(d["value"], 1), "it/s", round(d["ms_per_step"], 4), "tail", round(d["config"]["tail_ms_per_step"], 4), "scalar", round(d["config"]["tail_scalar_part_ms_per_step"], 4))
PY
