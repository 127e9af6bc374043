#!/usr/bin/env python
"""Per-kernel SASS summary of the in-tree library (runs on the CPU: cuobjdump only).

    python profiles/recipes/sass_summary.py stepth_b200/libstepth_cuda.so > profiles/r2_sass_summary.txt

For each kernel: instructions, registers, static shared memory, spill bytes (from `cuobjdump -res-usage`), and the counts
of the mnemonics that matter for this path: UTMALDG (2-D tensor-map TMA), UBLKCP (1-D bulk TMA), SYNCS (mbarrier),
tensor-core ops (none expected: the path is compare/select), FFMA (must be 0 in the finish kernels: unfused f32 rounding),
SHFL / VOTE / REDUX (warp exchanges), ATOM/RED, LDG/STG/LDS/STS.
"""
import re
import subprocess
import sys
from collections import Counter, OrderedDict

so = sys.argv[1] if len(sys.argv) > 1 else "stepth_b200/libstepth_cuda.so"
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
res = subprocess.run(["cuobjdump", "-res-usage", so], capture_output=True, text=True).stdout
demangle = lambda s: subprocess.run(["cu++filt", s], capture_output=True, text=True).stdout.strip() or s

usage = {}
cur = None
for line in res.splitlines():
    m = re.search(r"Function (\S+):", line)
    if m:
        cur = m.group(1)
        continue
    if cur and "REG:" in line:
        usage[cur] = dict(re.findall(r"(\w+):(\d+)", line))
        cur = None

KEYS = ["UTMALDG", "UTMASTG", "UBLKCP", "SYNCS", "HMMA", "IMMA", "UTCMMA", "FFMA", "FMUL", "FADD", "SHFL", "VOTE", "REDUX",
        "ATOM", "RED", "LDG", "STG", "LDS", "STS", "LOP3", "IMAD", "ISETP", "BRA"]
kern = OrderedDict()
cur = None
archs = set(re.findall(r"arch = (sm_\w+)", sass))
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        kern[cur] = Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
    if m and cur:
        op = m.group(1)
        kern[cur]["_n"] += 1
        kern[cur][op] += 1

print(f"# {so}: architectures {sorted(archs)}, {len(kern)} kernels")
tot = Counter()
for k, c in kern.items():
    u = usage.get(k, {})
    name = demangle(k)
    depth = 0                                  # drop the trailing parameter list, keep template arguments
    for i in range(len(name) - 1, -1, -1):
        depth += (name[i] == ")") - (name[i] == "(")
        if depth == 0 and name[i] == "(":
            name = name[:i]
            break
    cols = " ".join(f"{key}={c[key]}" for key in KEYS if c[key])
    print(f"{name}\n    inst={c['_n']} REG={u.get('REG','?')} SHARED={u.get('SHARED','?')} STACK={u.get('STACK','?')} "
          f"LOCAL={u.get('LOCAL','?')}\n    {cols}")
    for key in KEYS:
        tot[key] += c[key]
print("# totals: " + " ".join(f"{key}={tot[key]}" for key in KEYS))
fin = {k: c for k, c in kern.items() if "finish" in k}
print("# FFMA in finish_* kernels: " + str(sum(c["FFMA"] for c in fin.values())) + f" (over {len(fin)} kernels)")
