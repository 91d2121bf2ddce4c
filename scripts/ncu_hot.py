"""Per-instruction execution counts + stall samples from an ncu report (SASS page).
usage: python scripts/ncu_hot.py file.ncu-rep [min_share]   -> prints contiguous hot regions"""
import csv, subprocess, sys, io
f = sys.argv[1]
out = subprocess.run(['ncu', '-i', f, '--page', 'source', '--csv', '--print-source', 'sass',
                      '--launch-count', '1'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[1]
ix = {n: i for i, n in enumerate(hdr)}
data = []
for r in rows[2:]:
    if len(r) != len(hdr):
        if data:
            break          # next kernel instance
        continue
    if r[1] == 'Source':
        continue
    data.append(r)
tot = sum(int(r[ix['Instructions Executed']]) for r in data)
tots = sum(int(r[ix['# Samples']]) for r in data)
print('total inst', tot, 'samples', tots)
for k, r in enumerate(data):
    n = int(r[ix['Instructions Executed']]); s = int(r[ix['# Samples']])
    print(f"{k:5d} {n:10d} {100*n/tot:5.2f}% smp {100*s/max(tots,1):5.2f}%  {r[ix['Source']].strip()}")
