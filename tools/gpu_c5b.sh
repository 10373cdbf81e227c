#!/bin/bash
# Batched path only: its parity tests, the C5 bench line per setting, the launch list.
OUT=gpurun_out/${1:-c5b}
mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "batched" > $OUT/pytest_batched.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/pytest_batched.log
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py --workload c5 --steps 20 --warmup 3 --no-cpu > $OUT/bench_c5_$name.json 2> $OUT/bench_c5_$name.err
  echo "c5 $name rc=$? $(python -c "import json; d=json.load(open('$OUT/bench_c5_$name.json')); print(round(d['value'],1), 'batch-it/s', round(d['roofline']['achieved'],2), 'TF')" 2>&1 | tail -1)"
}
run default PPLS_X=0
run waves4 PPLS_BACC_WAVES=4
run ws0 PPLS_BGEMM_WS=0
run unfused PPLS_BATCH_FUSED=0 PPLS_BGEMM_WS=0
CMD="python bench.py --workload c5 --steps 4 --warmup 2 --no-cpu"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/c5_launches.csv $CMD > $OUT/ncu_c5.log 2>&1
python tools/launch_times.py $OUT/c5_launches.csv | tail -7
