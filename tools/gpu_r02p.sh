#!/bin/bash
# session P (1 GPU): full ncu capture of the update kernel at C4
set -u
OUT=gpurun_out/r02p
mkdir -p $OUT
CMD="python bench.py --workload c4 --steps 4 --warmup 3 --no-cpu --no-e2e --no-others"
$CMD > $OUT/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:em_update_grid -s 3 -c 1 -f -o $OUT/k2_c4_full $CMD > $OUT/ncu_full.log 2>&1
echo "ncu full rc=$?"
