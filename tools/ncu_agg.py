"""Aggregates an ncu --csv launch list (gpu__time_duration.sum [+ dram bytes]) per kernel name."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if r and r[0] == 'ID'][0]
hdr = rows[hi]
ix = {h: i for i, h in enumerate(hdr)}
agg = collections.OrderedDict()
scale = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9, 'Tbyte': 1e12}
for r in rows[hi + 1:]:
    if len(r) < len(hdr):
        continue
    name = r[ix['Kernel Name']][:int(sys.argv[2]) if len(sys.argv) > 2 else 70]
    m = r[ix['Metric Name']]
    v = float(r[ix['Metric Value']].replace(',', ''))
    u = r[ix['Metric Unit']]
    a = agg.setdefault(name, {'n': 0, 't': 0.0, 'rd': 0.0, 'wr': 0.0})
    if m == 'gpu__time_duration.sum':
        a['n'] += 1
        a['t'] += v / 1e6 if u == 'ns' else (v / 1e3 if u == 'us' else (v if u == 'ms' else v * 1e3))
    elif m == 'dram__bytes_read.sum':
        a['rd'] += v * scale[u]
    elif m == 'dram__bytes_write.sum':
        a['wr'] += v * scale[u]
tot = sum(a['t'] for a in agg.values())
print(f"| kernel | launches | time ms | share | DRAM read MB | DRAM write MB |\n|---|---|---|---|---|---|")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1]['t']):
    print(f"| `{k}` | {a['n']} | {a['t']:.3f} | {100 * a['t'] / tot:.1f} % | {a['rd'] / 1e6:.1f} | {a['wr'] / 1e6:.1f} |")
print(f"| total | | {tot:.3f} | | | |")
