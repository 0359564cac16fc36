"""Markdown tables of the committed bench lines: python tools/bench_tables.py r02b  (reads profiles/<tag>_bench*.json)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r02b"


def load(name):
    path = os.path.join(ROOT, "profiles", name)
    if not os.path.exists(path):
        return None
    lines = [l for l in open(path).read().strip().splitlines() if l.startswith("{")]
    return json.loads(lines[-1]) if lines else None


runs = {1: load(f"{tag}_bench.json")}
for n in (2, 4, 8):
    runs[n] = load(f"{tag}_bench_n{n}.json")
ref = load(f"{tag}_bench_reference.json")
runs = {n: d for n, d in runs.items() if d}
g = lambda x: f"{x / 1e9:.1f}"
out = []
out.append("| GPUs | device-resident, weak by time (10⁹ cells/s; ms/step) | × N=1 | end to end, host buffers (10⁹ cells/s; ms/step) | H2D all ranks / bare-copy ceiling (GB/s) | of ceiling | strong: one timestep over N ranks (ms; kernel ms busiest rank; exchange ms — the fastest of the measured exchange paths, §7) | × N=1 |")
out.append("|---|---|---|---|---|---|---|---|")
b = runs.get(1)
for n, d in sorted(runs.items()):
    e, s = d["e2e"], d.get("strong")
    st = "—"
    sx = "—"
    if s:
        c = s["collective"]
        st = f"{s['ms_per_step']:.2f}; {s['kernel_ms']:.2f}; {c['ms']:.2f}" + (" (copy only)" if n == 1 else "")
        if b and b.get("strong"):
            sx = f"{b['strong']['ms_per_step'] / s['ms_per_step']:.2f}"
    out.append(f"| {n} | {g(d['value'])}; {d['ms_per_step']:.2f} | {d['value'] / b['value']:.2f} | {g(e['value'])}; {e['ms_per_step']:.1f} | "
               f"{e.get('h2d_gbs_all_ranks', 0):.0f} / {e.get('h2d_ceiling_gbs', 0):.0f} | {e.get('frac_of_h2d_ceiling', 0):.2f} | {st} | {sx} |")
out.append("")
out.append("| GPUs | configs[1] model-level, 20 variables (10⁹ cells/s; % roofline) | configs[3] median + upsample sweep (10⁹ cells/s; % roofline) | configs[4] C3072 → C384 by (tile, band) (10⁹ cells/s; ms; kernel ms; exposed exchange ms) |")
out.append("|---|---|---|---|")
for n, d in sorted(runs.items()):
    c = d.get("configs", {})
    def cell(k, extra=False):
        v = c.get(k)
        if not v:
            return "—"
        t = f"{g(v['value'])}; {100 * v['roofline']['frac']:.0f} %"
        if extra:
            col = v.get("collective") or {}
            t = f"{g(v['value'])}; {v['ms_per_step']:.0f}; {v.get('kernel_ms', 0):.0f}; {col.get('ms', 0) if col else 0:.1f}"
        return t
    out.append(f"| {n} | {cell('configs[1]')} | {cell('configs[3]')} | {cell('configs[4]', True)} |")
out.append("")
if b:
    r = b["roofline"]
    out.append(f"Dominant kernel at N = 1: {r['ms']:.2f} ms = {r['achieved']:.0f} GB/s = {100 * r['frac']:.1f} % of {r['peak']:.0f} GB/s"
               + (f"; dram traffic {r['traffic'] / 1e9:.2f} GB = {r['traffic'] / r['algorithmic_bytes']:.2f} × algorithmic" if r.get("traffic") else "") + ".")
    cb = b.get("cpu_baseline")
    if cb:
        out.append(f"CPU port ({cb['kind']}, {cb['cores']} cores, {cb['sample']}): {cb['value'] / 1e6:.1f} × 10⁶ cells/s → "
                   f"{b['e2e']['value'] / cb['value']:.0f} × end to end, {b['value'] / cb['value']:.0f} × device-resident.")
if ref:
    out.append(f"Reference arm (`--impl reference`): {ref['value'] / 1e6:.1f} × 10⁶ cells/s, {ref['ms_per_step'] / 1e3:.1f} s per tile-sample step.")
print("\n".join(out))
