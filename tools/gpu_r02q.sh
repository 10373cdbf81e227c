#!/bin/bash
set -u
OUT=gpurun_out/r02q
mkdir -p $OUT
for CFG in "1 0" "1 6" "0 0"; do
  set -- $CFG
  PPLS_BPROJ_PP=$1 PPLS_BGEMM_DBG=$2 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"bprojmu|baccum" -c 8 --csv --log-file $OUT/pp_$1_$2.csv python tools/dbg_c5.py > /dev/null 2>&1
  python - <<PY
import csv, collections
try:
    rows=[r for r in csv.reader(open("$OUT/pp_$1_$2.csv")) if len(r)>10 and r[0].isdigit()]
    agg=collections.defaultdict(list)
    for r in rows: agg[r[4].split('(')[0][:30]].append(float(r[-1].replace(',','')))
    print("pp/dbg $1 $2:", {k: [round(x/1e3,1) for x in v[-3:]] for k,v in agg.items()})
except Exception as e:
    print("pp $1 $2: no data", e)
PY
done
