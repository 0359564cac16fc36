"""Stall / instruction summary of the streaming kernel by role (geometry warps, field warps): python tools/ncu_regions.py rep"""
import csv, collections, subprocess, sys, io
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[1]
ia = hdr.index('Address'); isrc = hdr.index('Source'); iin = hdr.index('Instructions Executed'); ismp = hdr.index('# Samples')
st = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
data = []
for r in rows[2:]:
    try: data.append((int(r[ia], 16), r))
    except ValueError: pass
base = data[0][0]
cs = [a - base for a, r in data if 'TRY_ALLOC' in r[isrc]][0]
exits = [a - base for a, r in data if r[isrc].strip().endswith('EXIT') and a - base > cs]
ce, oe = exits[0], exits[-1] + 16
print("field-warp loop: %d instructions" % ((ce - cs) // 16))
for name, lo, hi in (("geometry", 0, cs), ("field loop", cs, ce), ("output", ce, oe), ("out-of-line", oe, 1 << 30)):
    agg = collections.Counter(); tot = 0; ex = 0
    for a, r in data:
        if lo <= a - base < hi:
            for i in st:
                if r[i].isdigit(): agg[hdr[i]] += int(r[i])
            tot += int(r[ismp]); ex += int(r[iin])
    print(f"{name:12s} samples {tot:7d} inst {ex/1e6:8.1f}M", {k.replace('stall_', ''): round(100 * v / max(tot, 1), 1) for k, v in agg.most_common(9)})
