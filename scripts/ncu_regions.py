"""Groups the per-instruction table of scripts/ncu_hot.py into straight-line regions
(equal execution count) and prints each region's share of executed warp-instructions
and of stall samples.  usage: python scripts/ncu_hot.py rep > hot.txt; python scripts/ncu_regions.py hot.txt"""
import re, sys
pat = re.compile(r'\s*(\d+)\s+(\d+)\s+([\d.]+)% smp\s+([\d.]+)%\s+(.*)')
lines = open(sys.argv[1]).read().splitlines()
tot = int(lines[0].split()[2])
groups = []
for l in lines[1:]:
    m = pat.match(l)
    if not m:
        continue
    k, n, smp, src = int(m.group(1)), int(m.group(2)), float(m.group(4)), m.group(5)
    if groups and groups[-1][2] == n:
        groups[-1][1] = k; groups[-1][3] += 1; groups[-1][4] += smp
    else:
        groups.append([k, k, n, 1, smp, src])
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.7
for g in groups:
    share = g[2] * g[3] / tot * 100
    if share >= thr or g[4] >= 2 * thr:
        print(f"{g[0]:5d}-{g[1]:5d} count={g[2]:9d} ninstr={g[3]:4d} inst={share:5.1f}% stall_samples={g[4]:5.1f}%  {g[5][:48]}")
