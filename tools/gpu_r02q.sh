#!/bin/bash
set -u
OUT=gpurun_out/r02q
mkdir -p $OUT
for DBG in 0 7 8 5; do
  PPLS_BGEMM_DBG=$DBG ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"bprojmu|baccum" -c 8 --csv --log-file $OUT/st_$DBG.csv python tools/dbg_c5.py > /dev/null 2>&1
  python - <<PY
import csv, collections
try:
    rows=[r for r in csv.reader(open("$OUT/st_$DBG.csv")) if len(r)>10 and r[0].isdigit()]
    agg=collections.defaultdict(list)
    for r in rows: agg[r[4].split('(')[0][:30]].append(float(r[-1].replace(',','')))
    print("dbg $DBG:", {k: [round(x/1e3,1) for x in v[-3:]] for k,v in agg.items()})
except Exception as e:
    print("dbg $DBG: no data", e)
PY
done
