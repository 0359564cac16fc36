"""Opcode histogram of a SASS address range (static instruction budget of a loop body, no GPU needed).

usage: python tools/sass_hist.py <object-or-so> <function-substring> [lo hi]   (hex addresses)
Without a range: lists branch / barrier / call instructions so that the loop bounds can be read off.
"""
import collections
import re
import subprocess
import sys


def sass(path, fun):
    names = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True).stdout
    out, keep = [], False
    for line in names.splitlines():
        if "Function :" in line:
            keep = fun in line
            if keep:
                out.append(line)
            continue
        if keep:
            out.append(line)
    return out


INSTR = re.compile(r"/\*([0-9a-f]{4,5})\*/\s+(.*?);")


def main():
    path, fun = sys.argv[1], sys.argv[2]
    lines = sass(path, fun)
    ins = []
    for line in lines:
        m = INSTR.search(line)
        if m:
            ins.append((int(m.group(1), 16), m.group(2).strip()))
    if len(sys.argv) < 5:
        for a, t in ins:
            if re.search(r"\b(BRA|BAR|SYNCS|CALL|RET|EXIT|WARPSYNC|BSSY|BSYNC|UTMALDG)\b", t):
                print(f"{a:05x}  {t}")
        print(len(ins), "instructions")
        return
    lo, hi = int(sys.argv[3], 16), int(sys.argv[4], 16)
    hist = collections.Counter()
    n = 0
    for a, t in ins:
        if lo <= a < hi:
            t = re.sub(r"^@!?U?P\w+\s+", "", t)
            op = t.split()[0]
            base = op.split(".")[0]
            if base in ("LDS", "STS", "LDG", "STG"):
                base = ".".join(op.split(".")[:1] + [p for p in op.split(".")[1:] if p in ("64", "128")])
            hist[base] += 1
            n += 1
    for k, v in hist.most_common():
        print(f"{v:5d}  {k}")
    print(f"{n:5d}  TOTAL  [{lo:x}, {hi:x})")


if __name__ == "__main__":
    main()
