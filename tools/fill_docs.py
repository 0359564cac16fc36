"""Fill the @...@ markers of DESIGN.md / BASELINE.md from profiles/<tag>_* (one-off after the profile pass)."""
import json, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r02b"
P = lambda n: os.path.join(ROOT, "profiles", n)
bench = json.loads([l for l in open(P(f"{tag}_bench.json")).read().splitlines() if l.startswith("{")][-1])
r = bench["roofline"]
ncu = {}
for line in open(P(f"{tag}_ncu_plevel.txt")):
    if line.startswith("#") or not line.strip():
        continue
    parts = line.split()
    ncu[parts[0]] = parts[1]
inst = float(ncu["smsp__inst_executed.sum"]) / 1e9
issue = float(ncu["smsp__issue_active.avg.pct_of_peak_sustained_active"])
ncu_ms = float(ncu["gpu__time_duration.sum"])
table = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "bench_tables.py"), tag], capture_output=True, text=True).stdout.strip()
rows = ["| kernel / case | bytes | time | GB/s | % of 6551 GB/s |", "|---|---|---|---|---|"]
for line in open(P(f"{tag}_kernel_rooflines.txt")):
    m = re.match(r"(.*?)\s+([\d.]+) MB\s+([\d.]+) us\s+(\d+) GB/s\s+([\d.]+) % of \d+(.*)", line.rstrip())
    if m:
        rows.append(f"| {m.group(1).strip()}{m.group(6).strip() and ' ' + m.group(6).strip()} | {m.group(2)} MB | {m.group(3)} µs | {m.group(4)} | {m.group(5)} % |")
sub = {
    "@PAIR_MS@": f"{r['ms']:.2f}", "@PAIR_GBS@": f"{r['achieved']:.0f}", "@PAIR_FRAC@": f"{100 * r['frac']:.1f}",
    "@PAIR_TRAFFIC@": f"{r['traffic'] / 1e9:.2f}" if r.get("traffic") else "n/a",
    "@PAIR_TRATIO@": f"{r['traffic'] / r['algorithmic_bytes']:.2f}" if r.get("traffic") else "n/a",
    "@PAIR_INST@": f"{inst:.2f}", "@PAIR_ISSUE@": f"{issue:.0f}", "@PAIR_NCU_MS@": f"{ncu_ms:.2f}",
    "@BENCH_TABLE@": table, "@KERNEL_TABLE@": "\n".join(rows),
}
for name in ("DESIGN.md", "BASELINE.md"):
    path = os.path.join(ROOT, name)
    s = open(path).read()
    for k, v in sub.items():
        s = s.replace(k, v)
    open(path, "w").write(s)
    left = re.findall(r"@[A-Z_]+@", s)
    print(name, "markers left:", left)
