"""Summarises an ncu `--metrics gpu__time_duration.sum --csv` launch list per kernel.
    python tools/launch_summary.py profiles/r01h_launches.csv > profiles/r01h_launches_summary.txt"""
import csv
import re
import sys
from collections import defaultdict

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
hdr = rows[0]
ki, vi, ui = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
tot = defaultdict(float)
cnt = defaultdict(int)
for r in rows[1:]:
    name = re.sub(r'\(.*', '', r[ki])
    name = re.sub(r'^void ', '', name)
    if name.startswith('cub::') or 'cub::' in name:
        name = 'cub::' + re.sub(r'<.*', '', name.split('cub::')[-1].split('::')[-1])
    v = float(r[vi].replace(',', ''))
    v *= {'ns': 1e-3, 'us': 1.0, 'ms': 1e3, 'msecond': 1e3, 'usecond': 1.0, 'nsecond': 1e-3}.get(r[ui], 1.0)
    tot[name] += v
    cnt[name] += 1
T = sum(tot.values())
print('total_us %.1f launches %d' % (T, sum(cnt.values())))
for k in sorted(tot, key=lambda k: -tot[k]):
    print('%10.1f us %5.1f%% x%-4d %s' % (tot[k], 100 * tot[k] / T, cnt[k], k))
