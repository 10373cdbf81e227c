#!/bin/bash
# Round 2, session A (1 GPU): parity tests incl. the oracle/_ref pins, smoke, the default bench line (C3 + others).
set -u
OUT=gpurun_out/r02a
mkdir -p $OUT
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm,clocks.max.mem,power.limit --format=csv > $OUT/gpu.csv 2>&1
free -g > $OUT/host_mem.txt; nproc >> $OUT/host_mem.txt
ls -la oracle/_ref >> $OUT/host_mem.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log
tail -5 $OUT/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $OUT/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/smoke.log
SECONDS=0
timeout 900 python bench.py --steps 20 --warmup 3 > $OUT/bench_default.json 2> $OUT/bench_default.err; echo "bench rc=$? in ${SECONDS}s"
tail -c 3000 $OUT/bench_default.json; tail -5 $OUT/bench_default.err
SECONDS=0
timeout 600 python bench.py --impl reference --steps 20 --warmup 3 > $OUT/bench_reference.json 2> $OUT/bench_reference.err; echo "reference rc=$? in ${SECONDS}s"
tail -c 1200 $OUT/bench_reference.json; tail -3 $OUT/bench_reference.err
