"""Print the last step of an ncu gpu__time_duration launch list: python tools/parse_launches.py file.csv nsteps"""
import csv, sys
f = sys.argv[1]; nsteps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
lines = [l for l in open(f) if not l.startswith('==')]
rows = list(csv.DictReader(lines))
n = len(rows) // nsteps
tot = 0
for row in rows[(nsteps - 1) * n:]:
    t = float(row['Metric Value'].replace(',', '')); u = row['Metric Unit']
    t = t / 1000 if u == 'ns' else (t * 1000 if u == 'ms' else t)
    tot += t
    print(f"{row['ID']:>5} {row['Kernel Name'][:70]:70s} {row['Grid Size']:>14} {t:9.1f} us")
print('total us', round(tot, 1))
