"""Sum an ncu --csv launch list (gpu__time_duration.sum) by kernel name: python scripts/launch_summary.py launches.csv"""
import csv, sys, collections, re
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
hdr = rows[0]; ki = hdr.index("Kernel Name"); vi = hdr.index("Metric Value"); ui = hdr.index("Metric Unit")
tot = collections.OrderedDict()
for r in rows[1:]:
    v = float(r[vi].replace(",", "")); u = r[ui]
    v *= {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "nsecond": 1e-3, "ms": 1e3, "msecond": 1e3}.get(u, 1.0)
    name = re.sub(r"\(.*", "", r[ki])
    t = tot.setdefault(name, [0, 0.0]); t[0] += 1; t[1] += v
all_us = sum(t[1] for t in tot.values())
for k, (n, us) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print("%-60s launches %6d  total %10.1f us  avg %9.2f us  share %5.1f %%" % (k[:60], n, us, us / n, 100 * us / all_us))
