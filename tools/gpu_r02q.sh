#!/bin/bash
# session Q (1 GPU): uneven row splits of the batched accumulate kernel: parity, C5 by weight
set -u
OUT=gpurun_out/r02q
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -m gpu -q -x -k "batched" > $OUT/pytest_batched.log 2>&1; echo "pytest batched rc=$?"; tail -2 $OUT/pytest_batched.log | cut -c1-300
for W in 1.0 1.1 1.2 1.3; do
PPLS_BACC_MIXW=$W timeout 300 python bench.py --workload c5 --steps 30 --warmup 5 --no-cpu --no-e2e --no-others 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('c5 mixw $W:', round(d['value'],1), 'batch-it/s ms/step', round(d['ms_per_step'],4), 'frac', round(d['roofline']['frac'],3))"
done
