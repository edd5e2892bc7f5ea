"""Summarise an ncu launch list (csv from tools/gpu_ll.sh): python tools/ll_summary.py file.csv"""
import csv, collections, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10 and r[0].isdigit()]
agg = collections.OrderedDict()
for r in rows:
    k = (r[4][:56], r[8]); t = float(r[-1])
    a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += t
tot = sum(a[1] for a in agg.values())
for k, a in agg.items():
    print(f"{k[0]:58s} grid {k[1]:14s} n={a[0]:3d} avg {a[1]/a[0]/1e3:7.2f} us share {100*a[1]/tot:5.1f}%")
print(f"total {tot/1e3:.1f} us over {len(rows)} launches")
