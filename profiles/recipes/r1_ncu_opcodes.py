#!/usr/bin/env python
"""ncu_opcodes.py REPORT.ncu-rep -> executed warp instructions of the first kernel by SASS opcode"""
import csv, io, subprocess, sys, collections
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = next(r for r in rows if r and r[0] == "Address")
ii = hdr.index("Instructions Executed"); isrc = hdr.index("Source")
acc = collections.Counter(); tot = 0
for r in rows:
    if len(r) > ii and r[0].startswith("0x"):
        try: n = int(r[ii])
        except ValueError: continue
        op = r[isrc].strip().split()
        op = op[1] if op and op[0].startswith("@") else (op[0] if op else "?")
        acc[op.split(".")[0]] += n; tot += n
print(f"total {tot/1e6:.1f} M")
for op, n in acc.most_common(28): print(f"{n/tot*100:5.1f}%  {op}")
