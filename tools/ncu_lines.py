#!/usr/bin/env python
"""tools/ncu_lines.py REPORT.ncu-rep [top]: instructions executed and stall samples per source line (ncu source page)."""
import csv, re, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
lines, h, path = {}, None, "?"
for r in rows:
    if not r: continue
    if r[0] == "File Path": path = r[1]; continue
    if r[0] == "Line No": h = r; ie, ss = h.index("Instructions Executed") - len(h), h.index("# Samples") - len(h); continue
    if h is None or len(r) < len(h) or not r[0].isdigit(): continue
    try: n, s = int(r[ie] or 0), int(r[ss] or 0)
    except ValueError: continue
    lines[(path, int(r[0]))] = [r[1], n, s]
tot = sum(v[1] for v in lines.values()); tots = sum(v[2] for v in lines.values())
print(f"total warp instructions {tot}, samples {tots}")
for (f, ln), (src, n, s) in sorted(lines.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"{f.split('/')[-1][:10]:10s}{ln:5d} {100*n/tot:5.1f}% inst {100*s/max(tots,1):5.1f}% smp  {src.strip()[:100]}")
by = {}
for f in {k[0] for k in lines}:
    try: src_all = open(f).read().splitlines()
    except OSError: continue
    fn = "?"
    for i, l in enumerate(src_all, 1):
        if l.startswith("__device__") or l.startswith("__global__"):
            mm = re.search(r"(\w+)\s*\(", l); fn = mm.group(1) if mm else fn
        if (f, i) in lines:
            b = by.setdefault(fn, [0, 0]); b[0] += lines[(f, i)][1]; b[1] += lines[(f, i)][2]
print("-- by function")
for k, (n, s) in sorted(by.items(), key=lambda kv: -kv[1][0]): print(f"{k:24s} {100*n/tot:5.1f}% inst {100*s/max(tots,1):5.1f}% smp")
