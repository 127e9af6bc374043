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
def _ln(pat, src):
    for i, l in enumerate(src, 1):
        if pat in l: return i
    raise KeyError(pat)
_src = open("/root/repo/stepth_b200/csrc/search_hier.cuh").read().split("\n")
_marks = [("tile_min_key", "STEPTH_HD uint32_t tile_min_key"), ("HierQuery", "struct HierQuery {"), ("fine_test", "STEPTH_HD void fine_test"),
          ("coarse_window/side_flags", "STEPTH_HD unsigned long long coarse_window"), ("hier_rings (host)", "STEPTH_HD bool hier_rings("),
          ("HierArgs..", "struct HierArgs {"), ("query_of", "void query_of("), ("publish_result", "template <bool ONE_PAIR"),
          ("export_open", "void export_open("), ("scratch", "constexpr int WTASKS"), ("prune_sides", "uint32_t side_flags32("),
          ("emit_side", "unsigned long long emit_side("), ("walk: ring setup", "bool hier_rings_warp("),
          ("walk: pool listing", "// ---- pool the tile tests of the warp"), ("walk: pool tests", "const int n_pass ="),
          ("walk: ring end", "if (go && best != KEY_NONE"), ("near kernel body", "#ifndef STEPTH_NEAR_MINB"),
          ("far kernel body", "void __launch_bounds__(128, STEPTH_FAR_MINB)"), ("launch", "inline int32_t launch_search_hier")]
_pos = [(n, _ln(pat, _src)) for n, pat in _marks]
groups = [(n, a, (_pos[i + 1][1] - 1) if i + 1 < len(_pos) else 10**6) for i, (n, a) in enumerate(_pos)]
_unused = [("tile_min_key", 89, 136), ("fine_test", 159, 179), ("coarse_window/side_flags", 181, 195), ("query_of", 283, 297),
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
