#!/bin/bash
# session R (1 GPU): update kernel with the parallel Gram sum: parity and C4 / C2 timings
set -u
OUT=gpurun_out/r02r
mkdir -p $OUT
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_seq.py tests/test_gpu_ref_pin.py -m gpu -q -x > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $OUT/pytest.log | cut -c1-400
for WL in c4 c2; do
  timeout 300 python bench.py --workload $WL --steps 100 --warmup 10 --no-cpu --no-e2e --no-others 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$WL:', round(d['value'],1), 'it/s kernels', round(d['roofline']['kernel_ms'],4), 'tail', round(d['config']['tail_ms_per_step'],4))"
done
PPLS_FUSE_TAIL=0 timeout 300 python bench.py --workload c2 --steps 100 --warmup 10 --no-cpu --no-e2e --no-others 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('c2 three-kernel form:', round(d['value'],1), 'it/s kernels', round(d['roofline']['kernel_ms'],4), 'tail', round(d['config']['tail_ms_per_step'],4))"
