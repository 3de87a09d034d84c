"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list: the last step's kernels in order.
Usage: python tools/launches.py gpurun_out/X_launches.csv [kernels_per_step]"""
import csv, re, sys
lines = [l for l in open(sys.argv[1]) if not l.startswith('==')]
rows = list(csv.DictReader(lines))
def us(r):
    t = float(r['Metric Value'].replace(',', '')); u = r['Metric Unit']
    return t / 1e3 if u == 'ns' else t * 1e3 if u == 'ms' else t * 1e6 if u == 's' else t
names = [re.sub(r'^void ', '', re.sub(r'\(.*', '', r['Kernel Name'])).replace('<unnamed>::', '') for r in rows]
# the last step starts at the last k_preprocess launch
starts = [i for i, n in enumerate(names) if n.startswith('k_preprocess')]
s = starts[-1] if starts else 0
tot = 0
for i in range(s, len(rows)):
    if names[i].startswith('at::'): continue
    t = us(rows[i]); tot += t
    print(f"{i-s:3d} {names[i][:34]:34s} grid {rows[i]['Grid Size']:>16s} {t:9.1f} us")
print(f"total {tot:.1f} us")
