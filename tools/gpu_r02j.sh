#!/bin/bash
# session J (2 GPUs): full GPU suite after the staged scalar part, then C2/C3/C4 lines at 1 and 2 GPUs
set -u
OUT=gpurun_out/r02j
mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log
tail -8 $OUT/pytest_gpu.log | cut -c1-1500
for WL in c2 c4 c5; do
timeout 300 python bench.py --workload $WL --steps 100 --warmup 10 --no-e2e --no-others --no-cpu > $OUT/bench_${WL}_1gpu.json 2>$OUT/bench_${WL}_1gpu.err
python - <<PY
import json
d = json.load(open("$OUT/bench_${WL}_1gpu.json"))
c = d["config"]
print("$WL 1gpu", round(d["value"], 1), "it/s", round(d["ms_per_step"], 4), "tail", c.get("tail_ms_per_step"), "scalar", c.get("tail_scalar_part_ms_per_step"), "frac", round(d["roofline"]["frac"], 3))
PY
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --workload c2 --steps 100 --warmup 10 --no-e2e --no-others > $OUT/bench_c2_2gpu.json 2> $OUT/bench_c2_2gpu.err
python - <<PY
import json
for line in open("$OUT/bench_c2_2gpu.json"):
    if line.startswith("{"):
        d = json.loads(line)
        print("c2 2gpu", round(d["value"], 1), "it/s", round(d["ms_per_step"], 4), "ms/step tail", round(d["config"]["tail_ms_per_step"], 4), "scalar", round(d["config"]["tail_scalar_part_ms_per_step"], 4), d.get("parity", {}).get("pass"))
PY
