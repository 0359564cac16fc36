"""Per-source-line summary of an ncu report: python tools/ncu_lines.py gpurun_out/prof_stream.ncu-rep [N]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; N = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
cur = None; hdr = None; lines = {}
for r in rows:
    if not r: continue
    if r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
    if r[0] == 'Line No': hdr = r; continue
    if len(r) > 64 and r[2] == '-' and r[0].isdigit():
        try: lines[(cur, int(r[0]))] = (int(r[7]), int(r[6]), r[1], r)
        except ValueError: pass
names = [h.replace('stall_', '') for h in hdr[31:48]]
ti = sum(v[0] for v in lines.values()); ts = sum(v[1] for v in lines.values())
print(f"total inst {ti/1e6:.0f}M samples {ts}")
def show(items):
    for k, v in items:
        st = {n: int(v[3][31+i]) for i, n in enumerate(names) if v[3][31+i].isdigit() and int(v[3][31+i]) > 0.15 * max(v[1], 1)}
        print(f"{k[0][:18]:18s}{k[1]:5d} {v[0]/1e6:8.1f}M {v[1]:7d}  {v[2][:70]:70s} {st}")
print("-- by samples"); show(sorted(lines.items(), key=lambda kv: -kv[1][1])[:N])
print("-- by instructions"); show(sorted(lines.items(), key=lambda kv: -kv[1][0])[:N])
if len(sys.argv) > 3:
    print("-- regions")
    for spec in sys.argv[3:]:
        name, a, b = spec.split(':'); a = int(a); b = int(b)
        sel = [v for k, v in lines.items() if k[0].startswith('plevel_') and a <= k[1] <= b]
        agg = {}
        for v in sel:
            for i, n in enumerate(names):
                if v[3][31+i].isdigit(): agg[n] = agg.get(n, 0) + int(v[3][31+i])
        s = sum(v[1] for v in sel)
        print(f"{name:12s} inst {sum(v[0] for v in sel)/1e6:8.1f}M samples {s:7d}", {n: c for n, c in agg.items() if c > 0.05 * max(s, 1)})
