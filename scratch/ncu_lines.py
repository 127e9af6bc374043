#!/usr/bin/env python
"""ncu_lines.py REPORT KERNEL_REGEX [N] -> hottest source lines (instructions executed, samples) of a kernel"""
import csv, io, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 40
which = int(sys.argv[4]) if len(sys.argv) > 4 else 0
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", f"regex:{kern}"],
                     capture_output=True, text=True).stdout
# the output holds one block per (kernel launch, file); take launch `which`
blocks, cur, fname = [], None, None
launch = -1
seen_files = set()
for row in csv.reader(io.StringIO(raw)):
    if not row: continue
    if row[0] == "File Path":
        fname = row[1].split("/")[-1]
        continue
    if row[0] == "Function Name":
        continue
    if row[0] == "Line No":
        hdr = row
        key = fname
        if key in seen_files and cur is not None and False: pass
        cur = (fname, hdr, [])
        blocks.append(cur)
        continue
    if cur is not None:
        cur[2].append(row)
lines = {}
tot_i = tot_s = 0
for fname, hdr, rows in blocks:
    ii, si = hdr.index("Instructions Executed"), hdr.index("# Samples")
    for r in rows:
        if r[0] == "" or len(r) <= ii: continue   # SASS rows
        try: inst, samp = float(r[ii]), float(r[si])
        except ValueError: continue
        k = (fname, int(r[0]), r[1].strip()[:90])
        a = lines.setdefault(k, [0, 0]); a[0] += inst; a[1] += samp
        tot_i += inst; tot_s += samp
print(f"total warp-inst {tot_i/1e6:.1f}M samples {tot_s:.0f}")
for (f, ln, src), (i, s) in sorted(lines.items(), key=lambda kv: -kv[1][0])[:n]:
    print(f"{100*i/tot_i:5.1f}% inst {100*s/max(tot_s,1):5.1f}% smp  {f}:{ln:<4d} {src}")
