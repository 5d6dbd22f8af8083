"""SASS page of an ncu report grouped into runs of equal execution count: python tools/ncu_regions.py x.ncu-rep [min_pct]"""
import csv, subprocess, sys
out = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'source', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))[2:]
minp = float(sys.argv[2]) if len(sys.argv) > 2 else 0.4
groups = []
for i, r in enumerate(rows):
    n, s, t = int(r[5]), int(r[2]), float(r[8])
    if groups and groups[-1][2] == n:
        g = groups[-1]; g[1] = i; g[3] += s; g[4] += n; g[5] += t
    else:
        groups.append([i, i, n, s, n, t])
tot = sum(int(r[5]) for r in rows); samp = sum(int(r[2]) for r in rows)
print("warp instructions", tot, "samples", samp)
for g in groups:
    if g[4] < tot * minp / 100: continue
    cnt = g[1] - g[0] + 1
    print(f"lines {g[0]:4d}-{g[1]:4d} ({cnt:3d} instr) exec/instr {g[2]:9d} total {100*g[4]/tot:5.1f}% samples {100*g[3]/samp:5.1f}% lanes {g[5]/cnt:5.1f}  {rows[g[0]][1].strip()[:60]}")
