#!/bin/bash
# One GPU-box session: parity tests, the bench line, the ncu launch list and one full capture of K1.
# Usage (under gpurun): bash tools/gpu_round.sh [tag]
set -u
TAG=${1:-r01}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm,clocks.max.mem,power.limit --format=csv > $OUT/gpu.csv 2>&1
free -g > $OUT/host_mem.txt; nproc >> $OUT/host_mem.txt
python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $OUT/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/smoke.log
python bench.py --steps 20 --warmup 3 > $OUT/bench_c3.json 2> $OUT/bench_c3.err; echo "bench rc=$?"
tail -c 1500 $OUT/bench_c3.json
python bench.py --workload c2 --steps 50 --warmup 5 --no-cpu > $OUT/bench_c2.json 2> $OUT/bench_c2.err; echo "bench c2 rc=$?"
tail -c 800 $OUT/bench_c2.json
python bench.py --workload c4 --steps 50 --warmup 5 --no-cpu > $OUT/bench_c4.json 2> $OUT/bench_c4.err; echo "bench c4 rc=$?"
python bench.py --workload c5 --steps 20 --warmup 3 --no-cpu > $OUT/bench_c5.json 2> $OUT/bench_c5.err; echo "bench c5 rc=$?"
# ncu: launch list of a short bench run, then one full capture of the streaming kernel
CMD="python bench.py --workload c3 --rows 2000000 --steps 4 --warmup 3 --no-e2e --no-cpu"
$CMD > $OUT/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv $CMD > $OUT/ncu_launch.log 2>&1
echo "ncu launches rc=$?"
PCMD="python tools/prof_k1.py c3 2000000 3"
$PCMD > $OUT/plain_prof.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:em_stream -s 1 -c 2 -f -o $OUT/k1_full $PCMD > $OUT/ncu_full.log 2>&1
echo "ncu full rc=$?"
cat $OUT/plain_prof.log | tail -2
