"""Table of a scripts/ab_variants.sh log: python scripts/ab_summary.py gpurun_out/<dir>/ab.log [baseline variant]"""
import collections
import re
import sys
best = collections.OrderedDict(); crc = {}
v = None
for l in open(sys.argv[1]):
    m = re.match(r'== (\S+)', l)
    if m:
        v = m.group(1); continue
    m = re.match(r'(\S+)\s+(\d+) dispatches.*best of 3:\s+([\d.]+) Mrays/s.*rays (\d+)\s+image crc (\w+)', l)
    if m:
        k = (m.group(1), int(m.group(2)))
        best.setdefault(v, {}); best[v][k] = max(best[v].get(k, 0), float(m.group(3)))
        crc.setdefault(k, set()).add((m.group(4), m.group(5)))
base = sys.argv[2] if len(sys.argv) > 2 else next(iter(best))
jobs = list(best[base].keys())
print("%-10s" % "", *["%15s" % ("%s:%d" % k) for k in jobs])
for v, d in best.items():
    print("%-10s" % v, *["%8.0f %+5.1f%%" % (d.get(k, 0), 100 * (d.get(k, 0) / best[base][k] - 1)) for k in jobs])
print("distinct (rays, image crc) per job:", {"%s:%d" % k: len(s) for k, s in crc.items()})
