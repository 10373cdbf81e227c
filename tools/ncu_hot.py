"""Print the hot SASS instructions of an `ncu --page source --csv --print-source sass` dump."""
import csv, sys
path = sys.argv[1]; thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.004
rows = list(csv.reader(open(path)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]; data = [r for r in rows[hi + 1:] if len(r) == len(hdr) and r[0].startswith("0x")]
ix = {h: i for i, h in enumerate(hdr)}
tot = sum(int(r[ix['# Samples']]) for r in data)
print('total samples', tot, 'instructions', len(data))
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
agg = {h: sum(int(r[ix[h]]) for r in data) for h in stalls}
print({k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
for n, r in enumerate(data):
    s = int(r[ix['# Samples']])
    if s > tot * thr:
        top = sorted(((int(r[ix[h]]), h) for h in stalls), reverse=True)[:2]
        print(n, r[ix['Source']].strip()[:70].ljust(70), s, r[ix['Instructions Executed']], top)
