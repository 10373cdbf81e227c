"""Aggregate the sampled stalls of an `ncu --page source --csv --print-source sass` dump per SASS opcode."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]; data = [r for r in rows[hi + 1:] if len(r) == len(hdr) and r[0].startswith("0x")]
ix = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
agg = collections.defaultdict(lambda: collections.Counter())
for r in data:
    src = r[ix['Source']].strip().split()
    op = next((t for t in src if not t.startswith('@')), '?').split('.')[0]
    agg[op]['samples'] += int(r[ix['# Samples']])
    agg[op]['n'] += 1
    agg[op]['exec'] += int(r[ix['Instructions Executed']])
    for h in stalls:
        agg[op][h] += int(r[ix[h]])
tot = sum(v['samples'] for v in agg.values())
for op, c in sorted(agg.items(), key=lambda kv: -kv[1]['samples'])[:14]:
    top = sorted(((c[h], h) for h in stalls), reverse=True)[:4]
    print(f"{op:10s} static {c['n']:5d} exec {c['exec']:10d} samples {c['samples']:6d} ({100*c['samples']/tot:4.1f}%)  {top}")
