"""Summarise an ncu launch list (long-format csv of tools/gpu_ll*.sh: one row per launch and metric):
   python tools/ll_summary.py file.csv [launches_per_step]
Per kernel (name, grid): launches, average duration, DRAM bytes read / written per launch, share of the summed time."""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) >= 15 and r[0].isdigit()]
launch = collections.OrderedDict()   # id -> {name, grid, metrics}
for r in rows:
    L = launch.setdefault(int(r[0]), {"name": r[4], "grid": r[8], "m": {}})
    L["m"][r[12]] = float(r[14])
agg = collections.OrderedDict()
for L in launch.values():
    a = agg.setdefault((L["name"].split("(")[0][:44], L["grid"]), [0, 0.0, 0.0, 0.0])
    a[0] += 1
    a[1] += L["m"].get("gpu__time_duration.sum", 0.0)
    a[2] += L["m"].get("dram__bytes_read.sum", 0.0)
    a[3] += L["m"].get("dram__bytes_write.sum", 0.0)
tot = sum(a[1] for a in agg.values()) or 1.0
for k, a in agg.items():
    print(f"{k[0]:46s} grid {k[1]:14s} n={a[0]:3d} avg {a[1]/a[0]/1e3:7.2f} us  rd {a[2]/a[0]/1e6:8.2f} MB  wr {a[3]/a[0]/1e6:7.2f} MB  share {100*a[1]/tot:5.1f}%")
n = len(launch)
steps = n / int(sys.argv[2]) if len(sys.argv) > 2 else 1.0
print(f"total {tot/1e3:.1f} us over {n} launches; DRAM read {sum(a[2] for a in agg.values())/1e6:.1f} MB, written {sum(a[3] for a in agg.values())/1e6:.1f} MB"
      + (f"; per step ({steps:g} steps): {tot/1e3/steps:.1f} us, {(sum(a[2]+a[3] for a in agg.values()))/1e6/steps:.1f} MB" if len(sys.argv) > 2 else ""))
