"""Aggregate an ncu `--metrics gpu__time_duration.sum --csv` launch list by kernel name: launches, total us, share, average."""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 24
hdr = None
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    if len(r) > 5 and r[0] == 'ID':
        hdr = r
        continue
    if hdr and len(r) == len(hdr):
        d = dict(zip(hdr, r))
        try:
            v = float(d['Metric Value'].replace(',', ''))
        except ValueError:
            continue
        u = d['Metric Unit']
        v = v / 1000 if u == 'ns' else v * 1000 if u == 'ms' else v
        name = re.sub(r'\(.*', '', d['Kernel Name'])
        agg[name][0] += 1
        agg[name][1] += v
tot = sum(v[1] for v in agg.values())
for k, v in sorted(agg.items(), key=lambda x: -x[1][1])[:top]:
    print("%6d %9.1f us %5.1f%% avg %7.2f  %s" % (v[0], v[1], 100 * v[1] / tot, v[1] / v[0], k[:100]))
print("total %.1f us, %d launches" % (tot, sum(v[0] for v in agg.values())))
