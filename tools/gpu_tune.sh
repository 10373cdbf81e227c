#!/bin/bash
# Sweep K1 launch parameters at C3 and C2 widths (resident synthetic shard), long enough per
# configuration for the power cap to settle.
OUT=gpurun_out/${1:-tune}
mkdir -p $OUT
python tools/tune_k1.py c3 8000000 25 > $OUT/tune_c3.log 2>&1; cat $OUT/tune_c3.log
python tools/tune_k1.py c2 1000000 200 > $OUT/tune_c2.log 2>&1; cat $OUT/tune_c2.log
