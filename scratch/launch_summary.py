"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list: python scratch/launch_summary.py X.csv"""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = next(r for r in rows if "Kernel Name" in r)
ik, iv, im = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Name")
tot, cnt = collections.Counter(), collections.Counter()
for r in rows:
    if r is hdr or len(r) <= iv or r[im] != "gpu__time_duration.sum":
        continue
    name = re.sub(r"\(.*", "", r[ik])
    tot[name] += float(r[iv].replace(",", "")) / 1000.0
    cnt[name] += 1
total = sum(tot.values())
print(f"{'kernel':58s} {'launches':>8s} {'total us':>12s} {'avg us':>10s} {'share':>7s}")
for name, t in tot.most_common():
    print(f"{name[:58]:58s} {cnt[name]:8d} {t:12.1f} {t / cnt[name]:10.2f} {100 * t / total:6.1f}%")
