#!/bin/bash
# session L (1 GPU): single-pass wide kernel: stress against the two-pass form, parity, bench
set -u
OUT=gpurun_out/r02l
mkdir -p $OUT
REPS=60 timeout 300 python tools/dbg_wide3.py > $OUT/dbg_default.log 2>&1; echo "dbg default rc=$?"; tail -2 $OUT/dbg_default.log | cut -c1-300
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_seq.py -m gpu -q -k "wide" > $OUT/pytest_wide.log 2>&1; echo "pytest wide rc=$?"; tail -3 $OUT/pytest_wide.log | cut -c1-600
for i in 1 2; do
  timeout 300 python bench.py --workload c4 --steps 100 --warmup 10 --no-cpu --no-e2e --no-others 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('c4:', round(d['value'],1), 'it/s kernels', round(d['roofline']['kernel_ms'],4), 'tail', round(d['config']['tail_ms_per_step'],4), 'frac', round(d['roofline']['frac'],3))"
done
