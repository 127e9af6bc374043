#!/usr/bin/env python
"""ncu_lines2.py REPORT.ncu-rep [top] -> hottest CUDA source lines (warp instructions executed) of the first kernel"""
import csv, io, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
cur = None; rows = []; hdr = None
for row in csv.reader(io.StringIO(out)):
    if not row: continue
    if row[0] == "File Path": cur = row[1].split("/")[-1]; continue
    if row[0] == "Function Name": continue
    if row[0] == "Line No": hdr = row; continue
    if row[0] and row[0].isdigit():
        i_inst = hdr.index("Instructions Executed"); i_thr = hdr.index("Thread Instructions Executed"); i_s = hdr.index("# Samples")
        try: rows.append((cur, int(row[0]), row[1].strip()[:90], int(row[i_inst]), int(row[i_thr]), int(row[i_s])))
        except ValueError: pass
tot = sum(r[3] for r in rows); tots = sum(r[5] for r in rows)
print(f"total warp-inst {tot/1e6:.1f} M, samples {tots}")
for r in sorted(rows, key=lambda r: -r[3])[:top]:
    print(f"{r[3]/tot*100:5.1f}% inst {r[5]/max(tots,1)*100:5.1f}% smp  thr/inst {r[4]/max(r[3],1):4.1f}  {r[0]}:{r[1]:<4d} {r[2]}")
