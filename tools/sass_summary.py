#!/usr/bin/env python
"""Static SASS instruction histogram per kernel of the built library -> profiles/r02_sass_summary.txt.
Runs on the CPU container (cuobjdump needs no GPU):  python tools/sass_summary.py [lib] [out]"""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "openspades_b200/lib/libosb200.so"
out = sys.argv[2] if len(sys.argv) > 2 else "profiles/r02_sass_summary.txt"
WATCH = ["UBLKCP", "UTMALDG", "UTMASTG", "LDGSTS", "SYNCS", "LDG", "STG", "LDS", "STS", "ATOM", "ATOMG", "ATOMS", "RED", "BAR",
         "SHFL", "VOTE", "WARPSYNC", "REDUX", "MUFU", "FLO", "BREV", "F2I", "I2FP", "IMAD", "LOP3", "FADD", "FMUL", "FFMA"]
sass = subprocess.check_output(["cuobjdump", "-sass", lib], text=True)
res = subprocess.check_output(["cuobjdump", "-res-usage", lib], text=True)
regs = {}
cur = None
for line in res.splitlines():
    m = re.search(r"Function (\S+):", line)
    if m:
        cur = m.group(1)
    m = re.search(r"REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+)", line)
    if m and cur:
        regs[cur] = tuple(int(x) for x in m.groups())
kern = collections.OrderedDict()
cur = None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        kern[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
    if m and cur:
        kern[cur]["total"] += 1
        op = m.group(1)
        if op in WATCH:
            kern[cur][op] += 1


def demangle(n):
    try:
        return subprocess.check_output(["c++filt", "-p", n], text=True).strip()
    except Exception:
        return n


with open(out, "w") as f:
    f.write(f"# SASS instruction histogram per kernel of {lib} (sm_100a), `cuobjdump -sass`, static counts (tools/sass_summary.py).\n")
    f.write("# Bulk / async copy instructions (UBLKCP, UTMALDG, UTMASTG, LDGSTS, SYNCS = mbarrier): listed when present; the default build has\n"
            "# none (the bulk-copy variant of k_gather_fog is behind -DOSB_GATHER_TMA=1, DESIGN.md section 4.11).\n")
    f.write("# kernel | SASS instructions | registers / stack / static smem / local | watched opcodes\n")
    for k, c in sorted(kern.items(), key=lambda kv: demangle(kv[0])):
        if demangle(k).startswith("cub::") or "cub18CUB" in k:
            continue
        r = regs.get(k, (0, 0, 0, 0))
        ops = " ".join(f"{o}={c[o]}" for o in WATCH if c[o])
        f.write(f"{demangle(k):64s} {c['total']:6d} | regs {r[0]:3d} stack {r[1]:3d} smem {r[2]:6d} local {r[3]:3d} | {ops}\n")
print("wrote", out, len(kern), "functions")
