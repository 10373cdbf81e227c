#!/bin/bash
# C2 step-time breakdown on G GPUs with the peer exchange on / off.  Usage: bash tools/mg_c2.sh G tag
G=${1:-2}; OUT=gpurun_out/${2:-mg}; mkdir -p $OUT
show() { python -c "import json,sys;j=json.loads(open('$1').read().strip().splitlines()[-1]);print('$1', 'it/s %.1f step %.1f us K1 %.1f us tail %.1f us' % (j['value'], 1e3*j['ms_per_step'], 1e3*j['roofline']['kernel_ms'], 1e3*j['config']['tail_ms_per_step']), j['config']['exchange'])"; }
timeout -k 10 120 python bench.py --workload c2 --steps 100 --warmup 5 --no-e2e --no-cpu > $OUT/c2_1gpu.json 2> $OUT/c2_1gpu.err; show $OUT/c2_1gpu.json
for x in 1 0; do
  PPLS_XCHG=$x timeout -k 10 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 2952$x bench.py --gpus $G --workload c2 --steps 100 --warmup 5 --no-e2e > $OUT/c2_${G}gpu_x$x.json 2> $OUT/c2_${G}gpu_x$x.err; echo rc=$?
  show $OUT/c2_${G}gpu_x$x.json
done
