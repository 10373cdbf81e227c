#!/bin/bash
# C2 step-time breakdown on G GPUs with the peer exchange on / off.  Usage: bash tools/mg_c2.sh G tag [workload steps]
G=${1:-2}; OUT=gpurun_out/${2:-mg}; WL=${3:-c2}; ST=${4:-100}; mkdir -p $OUT
show() { python -c "import json,sys;j=json.loads(open('$1').read().strip().splitlines()[-1]);print('$1', 'it/s %.1f step %.1f us K1 %.1f us tail %.1f us' % (j['value'], 1e3*j['ms_per_step'], 1e3*j['roofline']['kernel_ms'], 1e3*j['config']['tail_ms_per_step']), j['config']['exchange'], j['config'].get('per_rank_ms_step_k1_tail'))"; }
if [ "$G" = 1 ]; then timeout -k 10 120 python bench.py --workload $WL --steps $ST --warmup 5 --no-e2e --no-cpu > $OUT/${WL}_1gpu.json 2> $OUT/${WL}_1gpu.err; show $OUT/${WL}_1gpu.json; exit; fi
for x in 1 0; do
  PPLS_XCHG=$x timeout -k 10 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 2952$x bench.py --gpus $G --workload $WL --steps $ST --warmup 5 --no-e2e > $OUT/${WL}_${G}gpu_x$x.json 2> $OUT/${WL}_${G}gpu_x$x.err; echo rc=$?
  show $OUT/${WL}_${G}gpu_x$x.json
done
