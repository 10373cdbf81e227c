#!/bin/bash
# Sweep K1 launch parameters at C3 and C2 widths (resident synthetic shard).
OUT=gpurun_out/${1:-tune}
mkdir -p $OUT
python tools/tune_k1.py c3 3000000 > $OUT/tune_c3.log 2>&1; tail -8 $OUT/tune_c3.log
python tools/tune_k1.py c2 1000000 > $OUT/tune_c2.log 2>&1; tail -8 $OUT/tune_c2.log
