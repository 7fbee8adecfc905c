"""Summarise an ncu launch list (--metrics gpu__time_duration.sum[,dram__bytes_read.sum,dram__bytes_write.sum] --csv).
usage: python scripts/launch_summary.py launches.csv [first_n_to_list] [--json out.json]"""
import csv, json, re, sys
fn = sys.argv[1]
lines = [l for l in open(fn) if not l.startswith('==')]
by = {}
for r in csv.DictReader(lines):
    e = by.setdefault(int(r['ID']), {'k': re.sub(r'\(.*', '', r['Kernel Name']).replace('void ', ''), 'grid': r.get('Grid Size')})
    v = float(r['Metric Value'].replace(',', ''))
    u = r['Metric Unit']
    if r['Metric Name'].startswith('gpu__time'):
        e['us'] = v / 1000 if u == 'ns' else v * (1e3 if u == 'ms' else 1)
    else:
        e[r['Metric Name']] = v * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}.get(u, 1)
rows = [by[i] for i in sorted(by)]
agg = {}
for e in rows:
    a = agg.setdefault(e['k'], {'us': 0.0, 'n': 0, 'rd': 0.0, 'wr': 0.0})
    a['us'] += e['us']; a['n'] += 1
    a['rd'] += e.get('dram__bytes_read.sum', 0.0); a['wr'] += e.get('dram__bytes_write.sum', 0.0)
tot = sum(e['us'] for e in rows)
for k, a in sorted(agg.items(), key=lambda x: -x[1]['us']):
    print('%-40s n=%3d total %9.1f us  share %.3f  dram rd %8.1f MB wr %8.1f MB  (%.0f GB/s)'
          % (k[:40], a['n'], a['us'], a['us'] / tot, a['rd'] / 1e6, a['wr'] / 1e6, (a['rd'] + a['wr']) / a['us'] / 1e3))
print('total us %.1f over %d launches' % (tot, len(rows)))
args = [a for a in sys.argv[2:] if not a.startswith('--')]
if args and args[0].isdigit():
    for i, e in enumerate(rows[:int(args[0])]):
        print(i, e['k'][:36], e['grid'], '%.1f' % e['us'])
if '--json' in sys.argv:
    out = sys.argv[sys.argv.index('--json') + 1]
    json.dump({'source': fn, 'launches': len(rows), 'total_us': tot,
               'kernels': {k: {'launches': a['n'], 'us': a['us'], 'share': a['us'] / tot, 'dram_read_bytes': a['rd'],
                               'dram_write_bytes': a['wr']} for k, a in agg.items()}}, open(out, 'w'), indent=1)
