"""Per-chunk averages of gpurun_out/pair_trace.txt (tools/pair_trace.py)"""
import re, sys
import numpy as np
rows = []
for l in open(sys.argv[1] if len(sys.argv) > 1 else 'gpurun_out/pair_trace.txt'):
    m = re.match(r'\s*(\d+)\s+(\d+) \|\s*(\d+)\s+(\d+)\s+(\d+)\s+(\d+) top\s+(\d+) tma\s+(\d+) \|\s*(\d+)\s+(\d+)\s+(\d+)\s+(\d+) \|\s*(\d+)\s+period (\d+)', l)
    if m: rows.append([int(x) for x in m.groups()])
    if l.startswith('total'): print(l.strip())
a = np.array(rows)
for c in range(28):
    sel = a[(a[:, 1] == c) & (a[:, 0] >= 28)]
    print(c, 'geom wait', int(sel[:, 3].mean()), 'work', int(sel[:, 4].mean()), 'top', int(sel[:, 6].mean()), 'tma', int(sel[:, 7].mean()), '| field wait_rec', int(sel[:, 8].mean()), 'work', int(sel[:, 10].mean()), 'period', int(sel[:, 13].mean()))
