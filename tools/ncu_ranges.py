"""Stall / instruction summary of a kernel by SASS address range (ranges read off tools/sass_hist.py):
python tools/ncu_ranges.py <rep> [--kernel substr] name:lo:hi ...   (hex offsets from the function start; hi = 0 -> end)"""
import collections
import csv
import io
import subprocess
import sys

args = sys.argv[1:]
rep = args.pop(0)
cmd = ["ncu", "-i", rep, "--page", "source", "--csv"]
if args and args[0] == "--kernel":
    cmd += ["-k", "regex:" + args[1]]
    args = args[2:]
out = subprocess.run(cmd, capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h = next(i for i, r in enumerate(rows) if "Address" in r)
hdr = rows[h]
ia, isrc = hdr.index("Address"), hdr.index("Source")
iin, ismp = hdr.index("Instructions Executed"), hdr.index("# Samples")
st = [i for i, c in enumerate(hdr) if c.startswith("stall_") and "Not Issued" not in c]
data = []
for r in rows[h + 1:]:
    try:
        data.append((int(r[ia], 16), r))
    except (ValueError, IndexError):
        pass
base = data[0][0]
ranges = [a.split(":") for a in args] or [["all", "0", "0"]]
total_s = sum(int(r[ismp]) for _, r in data)
total_i = sum(int(r[iin]) for _, r in data)
print(f"kernel: {len(data)} instructions, {total_s} samples, {total_i / 1e6:.1f}M warp instructions executed")
for name, lo, hi in ranges:
    lo, hi = int(lo, 16), int(hi, 16) or 1 << 30
    agg = collections.Counter()
    tot = ex = 0
    for a, r in data:
        if lo <= a - base < hi:
            for i in st:
                if r[i].isdigit():
                    agg[hdr[i]] += int(r[i])
            tot += int(r[ismp])
            ex += int(r[iin])
    top = {k.replace("stall_", ""): round(100 * v / max(tot, 1), 1) for k, v in agg.most_common(8)}
    print(f"{name:14s} samples {tot:7d} ({100 * tot / max(total_s, 1):4.1f} %) inst {ex / 1e6:8.1f}M", top)
