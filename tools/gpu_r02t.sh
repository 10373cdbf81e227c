#!/bin/bash
# session T (1 GPU): full ncu captures of the two batched GEMM kernels (C5)
set -u
OUT=gpurun_out/r02t
mkdir -p $OUT
python tools/dbg_c5.py > $OUT/plain.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"bprojmu|baccum" -s 8 -c 2 -f -o $OUT/c5_full python tools/dbg_c5.py > $OUT/ncu.log 2>&1
echo "ncu rc=$?"
