"""Summarise an ncu launch list (gpu__time_duration per launch) by kernel."""
import collections, csv, re, sys
path = sys.argv[1] if len(sys.argv) > 1 else 'gpurun_out/launches.csv'
lines = [l for l in open(path) if not l.startswith('==')]
agg = collections.defaultdict(lambda: [0, 0.0])
tot = 0
for row in csv.DictReader(lines):
    if row.get('Metric Name') != 'gpu__time_duration.sum':
        continue
    name = re.sub(r'\(.*', '', row['Kernel Name']).replace('void ', '').replace('gsb::', '')
    v = float(row['Metric Value'].replace(',', ''))
    u = row['Metric Unit']
    v = v / 1e3 if u == 'ns' else (v * 1e3 if u == 'ms' else v)
    agg[name][0] += 1
    agg[name][1] += v
    tot += v
print('total %.1f us over %d launches' % (tot, sum(a[0] for a in agg.values())))
for k, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1])[:int(sys.argv[2]) if len(sys.argv) > 2 else 24]:
    print('%-58s n=%4d %10.1f us %5.1f%%  avg %8.1f' % (k[:58], c, t, 100 * t / tot, t / c))
