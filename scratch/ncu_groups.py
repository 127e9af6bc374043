#!/usr/bin/env python
"""ncu_groups.py REPORT.ncu-rep -> warp instructions of search_hier.cuh grouped by code region"""
import csv, io, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
cur = None; hdr = None; rows = []
for row in csv.reader(io.StringIO(out)):
    if not row: continue
    if row[0] == "File Path": cur = row[1].split("/")[-1]; continue
    if row[0] == "Line No": hdr = row; continue
    if row[0] and row[0].isdigit():
        try: rows.append((cur, int(row[0]), int(row[hdr.index("Instructions Executed")]), int(row[hdr.index("Thread Instructions Executed")]), int(row[hdr.index("# Samples")])))
        except ValueError: pass
groups = [("tile_min_key", 89, 136), ("fine_test", 159, 179), ("coarse_window/side_flags", 181, 195), ("query_of", 283, 297),
          ("publish_result", 299, 312), ("export_open", 314, 328), ("tile_smin+emit_side", 345, 368), ("walk: ring setup", 372, 423),
          ("walk: pool listing", 424, 437), ("walk: pool tests", 438, 457), ("walk: ring end", 458, 462), ("near kernel body", 464, 549), ("far kernel body", 551, 585)]
tot = sum(r[2] for r in rows); acc = {}
for f, ln, inst, thr, smp in rows:
    g = f
    if f == "search_hier.cuh":
        g = next((n for n, a, b in groups if a <= ln <= b), "search_hier other")
    a = acc.setdefault(g, [0, 0, 0]); a[0] += inst; a[1] += thr; a[2] += smp
ts = sum(a[2] for a in acc.values())
print(f"total {tot/1e6:.1f} M warp-inst")
for g, a in sorted(acc.items(), key=lambda kv: -kv[1][0]):
    print(f"{a[0]/tot*100:5.1f}% inst  {a[2]/ts*100:5.1f}% samples  lanes {a[1]/max(a[0],1):4.1f}  {g}")
