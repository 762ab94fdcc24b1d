"""Summarise an ncu --metrics gpu__time_duration.sum CSV: per-kernel time inside the last decode step."""
import collections, csv, sys
lines = [l for l in open(sys.argv[1]) if not l.startswith('==')]
rows = list(csv.DictReader(lines))
names = [x['Kernel Name'] for x in rows]
idx = [i for i, n in enumerate(names) if 'cfg_sample' in n]
a, b = idx[-2] + 1, idx[-1] + 1
step = rows[a:b]
agg = collections.OrderedDict()
for x in step:
    k = (x['Kernel Name'].split('(')[0][:64], x['Grid Size'], x['Block Size'])
    d = agg.setdefault(k, [0, 0.0]); d[0] += 1; d[1] += float(x['Metric Value'])
tot = sum(v[1] for v in agg.values())
for k, v in agg.items():
    print(f"{k[0]:66s} grid={k[1]:12s} blk={k[2]:12s} n={v[0]:3d} avg={v[1]/v[0]/1e3:7.2f}us share={v[1]/tot*100:5.1f}%")
print(f"total {tot/1e3:.1f} us over {len(step)} launches")
