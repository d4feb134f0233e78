"""Aggregates `ncu --page source --print-source cuda,sass --csv` by CUDA source line.
usage: srclines.py file.csv [topN]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
cur = None; hdr = None; agg = collections.OrderedDict(); stall = collections.Counter()
for r in rows:
    if not r: continue
    if r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
    if r[0] == 'Function Name': continue
    if r[0] == 'Line No': hdr = r; iE = hdr.index('Instructions Executed'); iS = hdr.index('# Samples'); continue
    if hdr is None or len(r) <= iE: continue
    if r[0].strip().isdigit() and r[iE].strip().isdigit():   # the per-line roll-up row
        key = (cur, int(r[0]), r[1].strip()[:100])
        a = agg.setdefault(key, [0, 0]); a[0] += int(r[iE]); a[1] += int(r[iS]) if r[iS].strip().isdigit() else 0
tot = sum(a[0] for a in agg.values()); ts = sum(a[1] for a in agg.values())
print('total warp-instructions', tot, 'samples', ts)
n = int(sys.argv[2]) if len(sys.argv) > 2 else 45
for (f, l, s), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:n]:
    print(f'{100*a[0]/tot:5.1f}% inst {100*a[1]/max(ts,1):5.1f}% smp {a[0]:11d} {f}:{l}: {s}')
