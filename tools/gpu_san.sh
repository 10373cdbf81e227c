#!/bin/bash
# compute-sanitizer, ONE tool per gpurun call: bash tools/gpu_san.sh memcheck|racecheck|synccheck [--multi]
set -u
TOOL=$1; shift
TAG=$TOOL${1:+_multi}
OUT=gpurun_out/san
mkdir -p $OUT
python tools/san_cases.py "$@" > $OUT/plain_$TAG.log 2>&1 || { echo "plain run failed"; tail -20 $OUT/plain_$TAG.log; exit 1; }
tail -3 $OUT/plain_$TAG.log
timeout 2400 compute-sanitizer --tool $TOOL --print-limit 20 python tools/san_cases.py "$@" > $OUT/$TAG.log 2>&1
echo "compute-sanitizer $TOOL rc=$?"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|========= (Error|Warning|Race|Invalid|Barrier)" $OUT/$TAG.log | head -30
tail -5 $OUT/$TAG.log
