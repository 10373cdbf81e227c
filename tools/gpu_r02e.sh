#!/bin/bash
# Round 2, session E (2 GPUs): the multi-GPU tests again (torchrun forms, multi-device context), C5 row-sharded.
set -u
OUT=gpurun_out/r02e
mkdir -p $OUT
timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "two_gpus or multi_device" > $OUT/pytest_mgpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_mgpu.log
tail -30 $OUT/pytest_mgpu.log | cut -c1-1800
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --workload c5 --steps 20 --warmup 3 --no-e2e --no-others --no-parity > $OUT/bench_c5_2gpu.json 2> $OUT/bench_c5_2gpu.err
echo "bench c5 2gpu rc=$?"; tail -c 1800 $OUT/bench_c5_2gpu.json; tail -3 $OUT/bench_c5_2gpu.err
