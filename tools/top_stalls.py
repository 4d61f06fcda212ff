#!/usr/bin/env python
"""Top stalled SASS instructions of an `ncu --page source --csv` dump, with their dominant reasons.
    python tools/top_stalls.py <source.csv> [n]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
for i, r in enumerate(rows):
    if 'Instructions Executed' in r and '# Samples' in r:
        hdr, start = r, i + 1
        break
isrc, ismp = hdr.index('Source'), hdr.index('# Samples')
data = rows[start:]
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
tot = sum(int(r[ismp] or 0) for r in data)
agg = {h: sum(int(r[hdr.index(h)] or 0) for r in data) for h in stalls}
print('samples', tot, {k[6:]: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:10]})
for k in sorted(range(len(data)), key=lambda k: -int(data[k][ismp] or 0))[:n]:
    r = data[k]
    why = sorted(((int(r[hdr.index(h)] or 0), h[6:]) for h in stalls), reverse=True)[:2]
    print('%5d %5s  %-72s %s' % (k, r[ismp], r[isrc].strip()[:72], why))
