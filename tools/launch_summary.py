"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name. usage: launch_summary.py file.csv [last_n_launches]"""
import csv, collections, sys
lines = [l for l in open(sys.argv[1]) if not l.startswith('==')]
rows = [r for r in csv.DictReader(lines) if r.get('Metric Name') == 'gpu__time_duration.sum']
if len(sys.argv) > 2:
    rows = rows[-int(sys.argv[2]):]
agg = collections.OrderedDict()
for row in rows:
    k = row['Kernel Name'].split('(')[0][:60]
    v = float(row['Metric Value'].replace(',', ''))
    u = row['Metric Unit']
    v = v / 1e3 if u in ('ns', 'nsecond') else v * 1e3 if u in ('ms', 'msecond') else v  # -> us
    a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
print(f"total {tot:.1f} us over {len(rows)} launches")
for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{a[1]:10.1f} us {a[0]:4d}x {100*a[1]/tot:5.1f}%  {k}")
