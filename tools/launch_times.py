"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import collections, csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10 and r[0].isdigit()]
agg = collections.OrderedDict()
for r in rows:
    agg.setdefault(r[4].split('(')[0][:64], []).append(float(r[-1]) / 1e3)
tot = sum(sum(v) for v in agg.values())
for k, v in agg.items():
    print(k.ljust(66), "n=%-4d avg %9.1f us  share %5.1f%%" % (len(v), sum(v) / len(v), 100 * sum(v) / tot))
