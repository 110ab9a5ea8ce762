"""Aggregate an `ncu --csv --metrics gpu__time_duration.sum` log by kernel name. usage: agg_launches.py file.csv [nsteps]"""
import csv, sys
from collections import defaultdict
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = rows[0]
ik, iv, ig = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Grid Size')
n = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
agg = defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    try:
        v = float(r[iv].replace(',', ''))
    except ValueError:
        continue
    agg[r[ik][:100]][0] += 1
    agg[r[ik][:100]][1] += v
tot = sum(v for _, v in agg.values())
print('total %.3f ms over %d launches (/%g steps = %.3f ms/step)' % (tot / 1e6, sum(c for c, _ in agg.values()), n, tot / 1e6 / n))
for k, (c, v) in sorted(agg.items(), key=lambda x: -x[1][1])[:30]:
    print('%9.3f ms %5.1f%%  x%-5d %s' % (v / 1e6 / n, 100 * v / tot, c, k))
