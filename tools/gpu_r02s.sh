#!/bin/bash
# session S (2 GPUs): the exchange inside the fused tail: breakdown, parity of the row-sharded forms, recovery
set -u
PPLS_TAIL_VERBOSE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29595 bench.py --gpus 2 --workload c2 --steps 100 --warmup 10 --no-cpu --no-e2e --no-others --no-parity 2> gpurun_out/x2.err | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('c2 2 GPUs:', round(d['value'],1), d['config']['tail_ms_per_step'])"
grep "exchange inside" gpurun_out/x2.err | tail -2 | cut -c1-300
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "two_gpus or multi_device or recovers" 2>&1 | tail -3
