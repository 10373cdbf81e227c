#!/bin/bash
# GPU session for the batched path (C5): parity tests, then the fused projection kernel and the cp.async depth of
# the accumulate GEMM against the separate kernels; then the C3 bench line with the full-shard e2e call.
# Usage (under gpurun): bash tools/gpu_c5.sh [tag]
set -u
TAG=${1:-c5}
OUT=gpurun_out/$TAG
mkdir -p $OUT
free -g > $OUT/host_mem.txt; nproc >> $OUT/host_mem.txt
cat /sys/fs/cgroup/memory.max /sys/fs/cgroup/memory.current >> $OUT/host_mem.txt 2>&1
timeout 600 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $OUT/pytest_gpu.log
for cfg in "0 2" "1 2" "1 3" "1 4" "0 3"; do
  set -- $cfg
  PPLS_BATCH_FUSED=$1 PPLS_BACC_STAGES=$2 timeout 300 python bench.py --workload c5 --steps 20 --warmup 3 --no-cpu \
    > $OUT/bench_c5_f$1_s$2.json 2> $OUT/bench_c5_f$1_s$2.err
  echo "c5 fused=$1 stages=$2 rc=$? $(python -c "import json,sys; d=json.load(open('$OUT/bench_c5_f$1_s$2.json')); print(round(d['value'],1), 'batch-it/s', round(d['roofline']['achieved'],2), 'TF')" 2>&1 | tail -1)"
done
CMD="python bench.py --workload c5 --steps 4 --warmup 2 --no-cpu"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/c5_launches.csv $CMD > $OUT/ncu_c5.log 2>&1
echo "ncu c5 rc=$?"
timeout 900 python bench.py --steps 20 --warmup 3 --no-cpu > $OUT/bench_c3.json 2> $OUT/bench_c3.err; echo "bench c3 rc=$?"
tail -c 1200 $OUT/bench_c3.json; tail -3 $OUT/bench_c3.err
