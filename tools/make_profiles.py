"""Summarise gpurun_out/ (launch list + full ncu capture of the streaming kernel + bench lines) into profiles/.
    python tools/make_profiles.py r01b          # tag of the files written
Only reads gpurun_out/; run after tools/gpu_round.sh came back."""
import collections, csv, io, json, os, shutil, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = os.path.join(ROOT, "gpurun_out"); P = os.path.join(ROOT, "profiles")
tag = sys.argv[1] if len(sys.argv) > 1 else "r01"

# ---- launch list -----------------------------------------------------------------------------
rows = [r for r in csv.reader(open(os.path.join(G, "launches.csv"))) if r and not r[0].startswith("==")]
hdr = rows[0]; ik = hdr.index("Kernel Name"); iv = hdr.index("Metric Value"); iu = hdr.index("Metric Unit")
lib = collections.OrderedDict(); other = collections.OrderedDict()
for r in rows[1:]:
    if len(r) <= iv: continue
    us = float(r[iv].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3}.get(r[iu], 1.0)
    name = r[ik]
    d = lib if ("cg::" in name or "stream::" in name or "pairk::" in name or "plevel_pair" in name) else other
    n, t = d.get(name, (0, 0.0)); d[name] = (n + 1, t + us)
tot = sum(t for _, t in lib.values())
with open(os.path.join(P, f"{tag}_launches_summary.txt"), "w") as f:
    f.write("# ncu --metrics gpu__time_duration.sum --clock-control none -c 400: python bench.py --steps 2 --warmup 3 --no-cpu --e2e-steps 0 --config 2\n")
    f.write("# Cold-cache, serialised launches: compare SHARES, not absolutes.  The capture holds torch's synthetic input\n")
    f.write("# generation first, then 5 steps (3 warm-up + 2 timed).\n")
    f.write(f"# library kernels: {sum(n for n, _ in lib.values())} launches, {tot/1e3:.3f} ms total\n")
    f.write(f"{'kernel':92s} {'launches':>8s} {'total_us':>12s} {'us_each':>10s} {'share_of_step':>14s}\n")
    for k, (n, t) in sorted(lib.items(), key=lambda kv: -kv[1][1]):
        f.write(f"{k[:92]:92s} {n:8d} {t:12.1f} {t/n:10.1f} {100*t/tot:13.2f}%\n")
    f.write("# torch kernels in the same capture (synthetic input generation, outside the steps):\n")
    for k, (n, t) in sorted(other.items(), key=lambda kv: -kv[1][1])[:8]:
        f.write(f"{k[:92]:92s} {n:8d} {t:12.1f}\n")

# ---- full capture ----------------------------------------------------------------------------
rep = os.path.join(G, "prof_stream.ncu-rep")
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(io.StringIO(raw))); h, u, v = r[0], r[1], r[2]
keep = ["dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__time_duration.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "launch__block_size", "launch__grid_size", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "lts__t_sectors_srcunit_tex_op_read.sum",
        "sm__cycles_elapsed.avg.per_second", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tma.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active"]
keep += [k for k in h if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("_per_issue_active.ratio")]
with open(os.path.join(P, f"{tag}_ncu_plevel.txt"), "w") as f:
    f.write("# ncu --set full --clock-control none --import-source on -k regex:plevel_pair_kernel -s 2 -c 1: python bench.py --steps 1 --warmup 3 --no-cpu --e2e-steps 0 --config 2\n")
    f.write(f"# kernel: {v[h.index('Kernel Name')]}, C384 budget timestep (26 fields, 6 tiles x 79 levels)\n")
    for k in sorted(set(keep)):
        if k in h: f.write(f"{k:100s} {v[h.index(k)]:>20s} {u[h.index(k)]}\n")
    # warp-state samples and executed instructions by role: source-line ranges of plevel_pair.cu (tools/ncu_lines.py)
    src = open(os.path.join(ROOT, "fv3net_b200", "csrc", "plevel_pair.cu")).read().splitlines()
    def line_of(marker, start=0):
        return next(i + 1 for i, l in enumerate(src) if i >= start and marker in l)
    marks = [("geometry_general", "void geometry_step("), ("geometry_interior", "void geometry_interior("),
             ("extras_pair", "u64 extras_pair("), ("record_loads", "struct PairRec {"), ("end_pair", "EndOut end_pair("),
             ("load_q", "u64 load_q("), ("hot_step", "void hot_step("), ("kernel_setup", "plevel_pair_kernel(const __grid_constant__"),
             ("column_walker", "column walker: lane = fine column"), ("geometry_prologue", "geometry warp gw: lane = fine column"),
             ("geometry_loop", "warp gw takes level sub of the stages"), ("field_loop", "field threads: work item ="),
             ("output", "coarse outputs: the four pairs"), ("host", "---- host side")]
    pos = [(n, line_of(m)) for n, m in marks]
    specs = [f"{n}:{a}:{pos[i + 1][1] - 1}" for i, (n, a) in enumerate(pos[:-1])]
    reg = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_lines.py"), rep, "12"] + specs + ["shared_helpers:1:272"],
                         capture_output=True, text=True).stdout
    f.write("# top source lines by samples / instructions, then warp-state samples and executed instructions by region (tools/ncu_lines.py;\n")
    f.write("# shared_helpers = plevel_common.cuh lines, counted for whichever role inlines them):\n")
    for line in reg.strip().splitlines(): f.write("#   " + line[:230] + "\n")
def val(k):
    x = float(v[h.index(k)].replace(",", "")); un = u[h.index(k)]
    return x * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}.get(un, 1.0)
import hashlib
khash = hashlib.sha256(open(os.path.join(ROOT, "fv3net_b200", "csrc", "plevel_pair.cu"), "rb").read()
                       + open(os.path.join(ROOT, "fv3net_b200", "csrc", "plevel_common.cuh"), "rb").read()).hexdigest()[:16]
json.dump({"kernel": v[h.index("Kernel Name")], "kernel_source_sha256_16": khash, "dram_bytes_read": val("dram__bytes_read.sum"),
           "dram_bytes_write": val("dram__bytes_write.sum"), "plevel_cells": 1817247744,
           "source": f"profiles/{tag}_ncu_plevel.txt (ncu --set full, one launch, C384 budget timestep)"},
          open(os.path.join(P, "traffic.json"), "w"), indent=1)
for src, dst in (("bench.json", f"{tag}_bench.json"), ("bench_reference.json", f"{tag}_bench_reference.json")):
    if os.path.exists(os.path.join(G, src)): shutil.copy(os.path.join(G, src), os.path.join(P, dst))
print(open(os.path.join(P, f"{tag}_launches_summary.txt")).read()[:1800])

# ---- the other kernels (tools/kernel_rooflines.py under ncu): average duration per kernel --------------
kn = os.path.join(G, "kernels_ncu.csv")
if os.path.exists(kn):
    rows = [r for r in csv.reader(open(kn)) if r and not r[0].startswith("==")]
    hdr = rows[0]; ik = hdr.index("Kernel Name"); iv = hdr.index("Metric Value"); iu = hdr.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        if len(r) <= iv or not ("cg::" in r[ik] or "stream::" in r[ik] or "pairk" in r[ik]): continue
        us = float(r[iv].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3}.get(r[iu], 1.0)
        n, t, lo = agg.get(r[ik], (0, 0.0, 1e30)); agg[r[ik]] = (n + 1, t + us, min(lo, us))
    with open(os.path.join(P, f"{tag}_kernels_ncu.txt"), "w") as f:
        f.write("# ncu --metrics gpu__time_duration.sum --clock-control none: python tools/kernel_rooflines.py\n")
        f.write("# (cold-cache, serialised launches; 23 launches per case: 3 warm-up + 20 timed)\n")
        f.write(f"{'kernel':100s} {'launches':>8s} {'avg_us':>10s} {'min_us':>10s}\n")
        for k, (n, t, lo) in agg.items():
            f.write(f"{k[:100]:100s} {n:8d} {t/n:10.1f} {lo:10.1f}\n")
    if os.path.exists(os.path.join(G, "kernel_rooflines.txt")):
        shutil.copy(os.path.join(G, "kernel_rooflines.txt"), os.path.join(P, f"{tag}_kernel_rooflines.txt"))

# ---- the streaming remapper's capture -----------------------------------------------------------------------
rep2 = os.path.join(G, "prof_mappm.ncu-rep")
if os.path.exists(rep2):
    raw = subprocess.run(["ncu", "-i", rep2, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(io.StringIO(raw))); h, u, v = r[0], r[1], r[2]
    keep2 = ["gpu__time_duration.sum", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
             "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "dram__bytes_read.sum",
             "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
             "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
             "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
             "smsp__thread_inst_executed_per_inst_executed.ratio"]
    keep2 += [k for k in h if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("_per_issue_active.ratio")]
    src_csv = subprocess.run(["ncu", "-i", rep2, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                             capture_output=True, text=True).stdout
    cur = None; L = {}
    for row in csv.reader(io.StringIO(src_csv)):
        if not row: continue
        if row[0] == "File Path": cur = row[1].split("/")[-1]; continue
        if len(row) > 8 and row[0].isdigit() and row[2] == "-":
            try: L[(cur, int(row[0]))] = (int(row[7]), int(row[6]), row[1])
            except ValueError: pass
    with open(os.path.join(P, f"{tag}_ncu_mappm_stream.txt"), "w") as f:
        f.write("# ncu --set full --clock-control none --import-source on -k regex:mappm_stream_kernel -s 3 -c 1: python tools/mappm_bench.py 2\n")
        f.write(f"# kernel: {v[h.index('Kernel Name')]}, one C384 field (6 x 79 x 384 x 384) to a fine-grid result, kord 1, iv 1\n")
        if os.path.exists(os.path.join(G, "mappm_bench.log")):
            for line in open(os.path.join(G, "mappm_bench.log")): f.write("# events, 10 launches: " + line)
        for k in sorted(set(keep2)):
            if k in h: f.write(f"{k:100s} {v[h.index(k)]:>20s} {u[h.index(k)]}\n")
        f.write("# source lines by executed warp instructions (file line  Minst  samples  source):\n")
        for k, x in sorted(L.items(), key=lambda kv: -kv[1][0])[:24]:
            f.write(f"#   {k[0][:18]:18s}{k[1]:5d} {x[0]/1e6:8.1f}M {x[1]:7d}  {x[2][:110]}\n")
