#!/usr/bin/env python
"""ncu_lines.py REPORT.ncu-rep KERNEL_REGEX [TOP] -> executed warp instructions of a kernel by CUDA source line
(needs a capture taken with --import-source on and a library built with -lineinfo)"""
import csv, io, subprocess, sys, collections
rep, kre = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 60
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", f"regex:{kre}"],
                     capture_output=True, text=True).stdout
cur, hdr = None, None
acc = collections.OrderedDict()
ops = collections.Counter()
for r in csv.reader(io.StringIO(out)):
    if not r:
        continue
    if r[0] in ("File Path", "File Name"):
        cur = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        hdr = r
        ii, it, isamp = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
        continue
    if not hdr or len(r) <= ii:
        continue
    try:
        n, t, s = int(r[ii]), int(r[it]), int(r[isamp] or 0)
    except ValueError:
        continue
    if r[0].isdigit():          # a CUDA source line: the totals of the SASS rows that follow it
        a = acc.setdefault((cur, int(r[0])), [0, 0, 0, r[1].strip()[:100]])
        a[0] += n; a[1] += t; a[2] += s
    elif r[2].startswith("0x"):  # one SASS instruction
        op = r[3].strip().split()
        op = op[1] if op and op[0].startswith("@") else (op[0] if op else "?")
        ops[op.split(".")[0]] += n
tot = sum(a[0] for a in acc.values()) or 1
tott = sum(a[1] for a in acc.values())
tots = sum(a[2] for a in acc.values()) or 1
print(f"total {tot/1e6:.1f} M warp instructions, {tott/tot:.1f} lanes on average, {tots} samples")
byfile = collections.Counter()
for (f, _), a in acc.items():
    byfile[f] += a[0]
print("by file:", {f: f"{n/tot*100:.1f}%" for f, n in byfile.most_common()})
print("by opcode:", ", ".join(f"{op} {n/tot*100:.1f}%" for op, n in ops.most_common(16)))
for (f, l), a in sorted(acc.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{a[0]/tot*100:5.2f}%  lanes {a[1]/max(a[0],1):4.1f}  stall-samples {a[2]/tots*100:4.1f}%  {f}:{l}  {a[3]}")
