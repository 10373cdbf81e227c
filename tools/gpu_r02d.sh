#!/bin/bash
# Round 2, session D (2 GPUs): multi-GPU tests (torchrun forms + the single-process multi-device context), bench --gpus 2.
set -u
OUT=gpurun_out/r02d
mkdir -p $OUT
nvidia-smi --query-gpu=index,name --format=csv > $OUT/gpus.csv
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log
tail -25 $OUT/pytest_gpu.log
SECONDS=0
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29610 bench.py --gpus 2 --steps 20 --warmup 3 > $OUT/bench_2gpu.json 2> $OUT/bench_2gpu.err
echo "bench 2gpu rc=$? in ${SECONDS}s"; tail -c 2500 $OUT/bench_2gpu.json; tail -5 $OUT/bench_2gpu.err
for F in 1 0; do
PPLS_FUSE_TAIL=$F timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --workload c2 --steps 50 --warmup 5 --no-e2e --no-others > $OUT/bench_c2_2gpu_fuse$F.json 2> $OUT/bench_c2_2gpu_fuse$F.err
echo "bench c2 2gpu fuse=$F rc=$?"; tail -c 1500 $OUT/bench_c2_2gpu_fuse$F.json; tail -3 $OUT/bench_c2_2gpu_fuse$F.err
done
