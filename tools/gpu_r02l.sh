#!/bin/bash
# session L (1 GPU): single-pass wide kernel: stress against the two-pass form, parity, bench
set -u
OUT=gpurun_out/r02l
mkdir -p $OUT
REPS=100 timeout 300 python tools/dbg_wide3.py > $OUT/dbg_default.log 2>&1; echo "dbg default rc=$?"; tail -3 $OUT/dbg_default.log | cut -c1-300
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_seq.py -m gpu -q -k "wide" > $OUT/pytest_wide.log 2>&1; echo "pytest wide rc=$?"; tail -5 $OUT/pytest_wide.log | cut -c1-600
for CFG in "1 4 4 1" "1 4 4 1" "0 3 4 1"; do
  set -- $CFG
  PPLS_WIDE_SINGLE=$1 PPLS_WIDE_LAG=$2 PPLS_WIDE_STAGES=$3 PPLS_WIDE_HINT=$4 timeout 300 python bench.py --workload c4 --steps 100 --warmup 10 --no-cpu --no-e2e --no-others 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('single/lag/stages/hint $CFG:', round(d['value'],1), 'it/s kernels', round(d['roofline']['kernel_ms'],4), 'tail', round(d['config']['tail_ms_per_step'],4), 'frac', round(d['roofline']['frac'],3))"
done
