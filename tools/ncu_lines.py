#!/usr/bin/env python
"""Per-source-line instruction counts from an ncu report: ncu_lines.py <rep> <kernel regex> [top N]"""
import csv, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "-k", f"regex:{kern}"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
cur_file = None; hdr = None; agg = []
for r in rows:
    if len(r) == 2 and r[0] == "File Path": cur_file = r[1].split("/")[-1]; continue
    if r and r[0] == "Line No": hdr = r; continue
    if hdr is None or len(r) < len(hdr): continue
    if r[0] not in ("", ) and r[0].isdigit():
        ie = hdr.index("Instructions Executed"); te = hdr.index("Thread Instructions Executed"); sm = hdr.index("# Samples")
        try: agg.append((int(r[ie]), int(r[te]), int(r[sm]), cur_file, int(r[0]), r[1].strip()[:110]))
        except ValueError: pass
tot = sum(a[0] for a in agg); tots = sum(a[2] for a in agg)
print(f"total warp-inst {tot}, samples {tots}")
key = (lambda a: a[2]) if "--by-samples" in sys.argv else (lambda a: a[0])
for a in sorted(agg, key=key, reverse=True)[:top]:
    print(f"{a[0]:>12} {100*a[0]/tot:5.1f}%  smp {100*a[2]/max(tots,1):5.1f}%  {a[3]}:{a[4]:<5} {a[5]}")
