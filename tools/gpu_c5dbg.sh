#!/bin/bash
# Timing experiment: the two GEMM kernels of the batched path with the copies or the DMMAs switched off.
OUT=gpurun_out/${1:-c5dbg}
mkdir -p $OUT
CMD="python tools/c5_dbg.py"
for d in 0 1 2; do
echo "dbg=$d"
PPLS_BGEMM_DBG=$d timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/c5_launches_d$d.csv $CMD > $OUT/ncu_c5_d$d.log 2>&1
python tools/launch_times.py $OUT/c5_launches_d$d.csv | grep "bprojmu\|baccum"
done
