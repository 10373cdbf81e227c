#!/bin/bash
# ncu --set full of the two GEMM kernels of the batched path (one launch each, after warm-up)
OUT=gpurun_out/${1:-ncu_c5}
mkdir -p $OUT
CMD="python tools/c5_dbg.py"
$CMD > $OUT/plain.log 2>&1 || { echo plain failed; tail -5 $OUT/plain.log; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"bprojmu|baccum" -s 4 -c 2 -f -o $OUT/c5_full $CMD > $OUT/ncu_full.log 2>&1
echo "ncu rc=$?"; tail -3 $OUT/ncu_full.log
