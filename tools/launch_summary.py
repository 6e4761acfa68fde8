"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list: per-kernel totals of the LAST step
(from the last launch whose name contains argv[2], default: whole file)."""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr, L = None, []
for r in rows:
    if r[0] == 'ID':
        hdr = r
        continue
    if hdr is None:
        continue
    name = r[hdr.index('Kernel Name')]
    v = float(r[hdr.index('Metric Value')].replace(',', ''))
    u = r[hdr.index('Metric Unit')]
    v = v / 1000 if u == 'ns' else (v * 1000 if u == 'ms' else v)
    L.append((name.split('(')[0][-64:], v, r[hdr.index('Grid Size')]))
if len(sys.argv) > 2:
    idx = [i for i, (n, _, _) in enumerate(L) if sys.argv[2] in n]
    L = L[idx[-1]:]
agg = collections.defaultdict(lambda: [0, 0.0])
for n, v, g in L:
    agg[n][0] += 1
    agg[n][1] += v
tot = sum(v[1] for v in agg.values())
print('total us %.1f launches %d' % (tot, len(L)))
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%-66s n=%4d total_us=%10.1f avg=%8.1f share=%.3f" % (k, v[0], v[1], v[1] / v[0], v[1] / tot))
print([(n[-30:], round(v), g) for n, v, g in L if v > 200])
