"""Summarise an ncu launch list (gpu__time_duration per launch) by kernel.
usage: launch_summary.py launches.csv [top_n] [--one-step]   (--one-step: only the launches from the first pack_weight_multi_kernel
up to the next one, i.e. exactly one training step of the U-Net bench)"""
import collections, csv, re, sys
args = [a for a in sys.argv[1:] if not a.startswith('--')]
one_step = '--one-step' in sys.argv
path = args[0] if args else 'gpurun_out/launches.csv'
top = int(args[1]) if len(args) > 1 else 24
lines = [l for l in open(path) if not l.startswith('==')]
rows = []
for row in csv.DictReader(lines):
    if row.get('Metric Name') != 'gpu__time_duration.sum':
        continue
    name = re.sub(r'\(.*', '', row['Kernel Name']).replace('void ', '').replace('gsb::', '')
    try:
        v = float(row['Metric Value'].replace(',', ''))
    except ValueError:
        continue
    u = row['Metric Unit']
    v = v / 1e3 if u == 'ns' else (v * 1e3 if u == 'ms' else v)
    rows.append((name, v))
if one_step:
    starts = [i for i, (n, _) in enumerate(rows) if n.startswith('pack_weight_multi_kernel')]
    if len(starts) >= 2:
        rows = rows[starts[0]:starts[1]]
agg = collections.defaultdict(lambda: [0, 0.0])
tot = 0.0
for name, v in rows:
    agg[name][0] += 1
    agg[name][1] += v
    tot += v
print('total %.1f us over %d launches' % (tot, len(rows)))
for k, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1])[:top]:
    print('%-58s n=%4d %10.1f us %5.1f%%  avg %8.1f' % (k[:58], c, t, 100 * t / max(tot, 1e-9), t / c))
