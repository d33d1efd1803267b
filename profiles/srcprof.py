"""Per-source-line instruction / stall-sample shares from
   ncu -i X.ncu-rep --page source --csv --print-source sass,cuda -k regex:KERNEL > dump.csv
usage: python profiles/srcprof.py dump.csv [min_pct]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
h = next(i for i, r in enumerate(rows) if 'Instructions Executed' in r)
hdr = rows[h]
ii = hdr.index('Instructions Executed'); si = hdr.index('# Samples'); ti = hdr.index('Thread Instructions Executed')
out = []
for r in rows[h + 1:]:
    if not r or not r[0].strip().isdigit():
        continue
    try:
        out.append((int(r[0]), int(r[ii]), int(r[si]), int(r[ti]), r[1]))
    except ValueError:
        continue
tot = sum(o[1] for o in out); stot = sum(o[2] for o in out)
print("total warp inst", tot, "stall samples", stot)
for ln, n, sm, tn, src in sorted(out):
    if n > thr / 100 * tot or sm > thr / 100 * stot:
        print(f"{ln:5d} {100*n/tot:5.1f}% inst {100*sm/stot:5.1f}% smp  act {tn/max(n,1):4.1f}  {src.strip()[:96]}")
