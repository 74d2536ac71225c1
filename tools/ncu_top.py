"""Top stalled SASS instructions of an .ncu-rep (source page): python tools/ncu_top.py file.ncu-rep [n]"""
import csv, subprocess, sys
rep = sys.argv[1]; n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
lines = txt.splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
rows = list(csv.DictReader(lines[start:]))
k = 'Warp Stall Sampling (All Samples)'
f = lambda x: float(x) if x and x.replace('.', '', 1).isdigit() else 0.0
tot = sum(f(r[k]) for r in rows)
for i, r in enumerate(rows): r['i'] = i
print("total samples", tot, "instructions", len(rows))
for r in sorted(rows, key=lambda r: -f(r[k]))[:n]:
    print("%6.0f %5.1f%%  #%5d exec %9s  %s" % (f(r[k]), 100 * f(r[k]) / tot, r['i'], r['Instructions Executed'], r['Source'].strip()[:100]))
det = subprocess.run(["ncu", "-i", rep, "--page", "details", "--csv"], capture_output=True, text=True).stdout
for r in csv.DictReader(det.splitlines()):
    if r['Metric Name'] in ('Duration', 'Issue Slots Busy', 'Executed Ipc Active', 'No Eligible', 'Active Warps Per Scheduler', 'Registers Per Thread', 'L2 Hit Rate', 'Mem Busy'):
        print(r['Metric Name'], r['Metric Value'], r['Metric Unit'])
