"""Static opcode counts of the library's kernels and the first hot step of the pair-thread kernel (no GPU needed):
    python tools/sass_excerpt.py r02b   -> profiles/r02b_sass_excerpt.txt"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r02b"
so = os.path.join(ROOT, "fv3net_b200", "lib", "libcoarsen_b200.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
fn, H = None, collections.OrderedDict()
pat = re.compile(r"/\*[0-9a-f]{4,5}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)")
for l in out.splitlines():
    m = re.search(r"Function : (\S+)", l)
    if m:
        fn = m.group(1); H[fn] = collections.Counter(); continue
    m = pat.search(l)
    if m and fn:
        H[fn][m.group(1)] += 1
dem = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()[:110]
keys = ["UTMALDG", "SYNCS", "USETMAXREG", "LDGSTS", "FFMA2", "FMUL2", "FADD2", "FMNMX3", "FMNMX", "FSET", "BAR", "MUFU", "LDS", "STS", "SHFL", "LDG", "STG"]
want = re.compile(r"plevel|mappm_stream|median_f8|mode_small_kernel<float, 8, 1>|wavg_f8_kernel<2>|wavg_f16_kernel<2>|upsample_kernel<float>|exposed_area_strip|multicast_store|replicated_outputs")
with open(os.path.join(ROOT, "profiles", f"{tag}_sass_excerpt.txt"), "w") as f:
    f.write("# cuobjdump -sass fv3net_b200/lib/libcoarsen_b200.so (nvcc 12.9, -gencode arch=compute_100a,code=sm_100a): static opcode counts per kernel,\n")
    f.write("# filtered to the Blackwell-specific or design-relevant mnemonics (prefix match: SYNCS = mbarrier ops, LDGSTS = cp.async,\n")
    f.write("# UTMALDG = TMA tensor load, USETMAXREG = setmaxnreg, FFMA2/FMUL2/FADD2 = packed f32x2, BAR = named barriers; multimem.st assembles to\n")
    f.write("# STG.E.128.STRONG.SYS on the multicast address).  No UTCMMA / TMEM: nothing on this path is a contraction.\n")
    f.write(f"{'kernel':112s} {'total':>6s} " + " ".join(f"{k:>6s}" for k in keys) + "\n")
    for fn, c in H.items():
        name = dem(fn)
        if sum(c.values()) < 50 or not want.search(name):
            continue
        row = [sum(v for k2, v in c.items() if k2.split(".")[0] == k) for k in keys]
        f.write(f"{name:112s} {sum(c.values()):6d} " + " ".join(f"{v:6d}" for v in row) + "\n")
    f.write("\n# plevel_pair_kernel<false>: first hot step of a stage (one column pair, level and field per thread), cuobjdump -sass of build/plevel_pair.o\n")
    sass = subprocess.run(["cuobjdump", "-sass", os.path.join(ROOT, "fv3net_b200", "build", "plevel_pair.o")], capture_output=True, text=True).stdout
    keep, n, seen_hot = False, 0, False
    for l in sass.splitlines():
        if "Function :" in l:
            keep = "plevel_pair_kernelILb0" in l; continue
        if not keep:
            continue
        m = re.search(r"/\*([0-9a-f]{4,5})\*/\s+(.*?);", l)
        if not m:
            continue
        a, t = int(m.group(1), 16), re.sub(" +", " ", m.group(2))
        if n == 0 and "LDS.128" in t and a > 0xd000:
            n = 1
        if n:
            f.write(f"  /*{a:05x}*/ {t}\n"); n += 1
            if n > 2 and "CALL.REL.NOINC" in t:      # the out-of-line call for further events ends the step
                break
print("written")
