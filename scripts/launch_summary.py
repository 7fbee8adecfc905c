"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list."""
import csv, re, sys
def load(fn):
    lines = [l for l in open(fn) if not l.startswith('==')]
    return [(re.sub(r'\(.*', '', r['Kernel Name']), r.get('Grid Size'), float(r['Metric Value'].replace(',', '')) / 1000)
            for r in csv.DictReader(lines)]
rows = load(sys.argv[1])
agg = {}
for k, g, v in rows:
    agg.setdefault(k, [0, 0]); agg[k][0] += v; agg[k][1] += 1
tot = sum(v for _, _, v in rows)
for k, (v, n) in sorted(agg.items(), key=lambda x: -x[1][0]):
    print('%-44s n=%3d total %9.1f us  share %.3f' % (k[:44], n, v, v / tot))
print('total us %.1f over %d launches' % (tot, len(rows)))
if len(sys.argv) > 2:
    for i, (k, g, v) in enumerate(rows[:int(sys.argv[2])]):
        print(i, k[:36], g, '%.1f' % v)
