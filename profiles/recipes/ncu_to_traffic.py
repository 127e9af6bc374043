#!/usr/bin/env python
"""ncu_to_traffic.py REPORT.ncu-rep KEY [PAIRS [SMS]] -> one entry of profiles/traffic.json (printed as JSON)

From one `ncu --set full --clock-control none` capture of ONE launch sequence of the search stage (tile_planes,
coarse_transpose, search_near, search_far):
  traffic   dram__bytes_read.sum + dram__bytes_write.sum summed over those kernels (bytes per launch sequence)
  issue     the instruction-issue roofline of the dominant kernel (the longest one):
            frac = smsp__inst_executed.sum / (SMs x 4 schedulers x sm cycles elapsed)  -- one warp instruction per
            scheduler per cycle is the ceiling; lanes_active = smsp__thread_inst_executed_per_inst_executed.ratio (of 32)
bench.py copies the entry into roofline.traffic / roofline.issue."""
import csv, io, json, subprocess, sys

rep, key = sys.argv[1], sys.argv[2]
pairs = int(sys.argv[3]) if len(sys.argv) > 3 else 16   # pairs in the captured launch sequence
sms = int(sys.argv[4]) if len(sys.argv) > 4 else 148
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}


def val(r, name, default=None):
    if name not in idx or r[idx[name]] in ("", "n/a"):
        return default
    v = float(r[idx[name]].replace(",", ""))
    u = units[idx[name]].lower()
    scale = {"kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "msecond": 1e-3, "usecond": 1e-6, "nsecond": 1e-9, "second": 1.0,
             "ms": 1e-3, "us": 1e-6, "ns": 1e-9}.get(u, 1.0)
    return v * scale


stage = ("tile_planes", "coarse_transpose", "search_near", "search_otile", "search_far")
seq, seen = [], set()
for r in rows[2:]:
    name = r[idx["Kernel Name"]]
    short = next((s for s in stage if s in name), None)
    if short is None or short in seen:
        continue       # the first launch of each stage kernel = one launch sequence
    seen.add(short)
    seq.append((short, name, r))
traffic = sum((val(r, "dram__bytes_read.sum", 0) + val(r, "dram__bytes_write.sum", 0)) for _, _, r in seq)
dom = max(seq, key=lambda t: val(t[2], "gpu__time_duration.sum", 0))
r = dom[2]
inst = val(r, "smsp__inst_executed.sum")
dur = val(r, "gpu__time_duration.sum")
cyc = val(r, "sm__cycles_elapsed.max") or val(r, "smsp__cycles_elapsed.max")
hz = val(r, "smsp__cycles_elapsed.avg.per_second") or val(r, "sm__cycles_elapsed.avg.per_second")
if cyc is None:
    cyc = dur * (hz or 1.965e9)
entry = {
    "traffic": int(traffic),
    "pairs_per_launch": pairs,
    "traffic_per_kernel": {s: int(val(rr, "dram__bytes_read.sum", 0) + val(rr, "dram__bytes_write.sum", 0)) for s, _, rr in seq},
    "duration_us_per_kernel": {s: round(val(rr, "gpu__time_duration.sum", 0) * 1e6, 1) for s, _, rr in seq},
    "issue": {
        "kernel": dom[1][:80], "inst_executed": int(inst), "duration_us": round(dur * 1e6, 1), "sm_cycles": int(cyc),
        "frac": round(inst / (sms * 4 * cyc), 4),
        "issue_active_pct": val(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
        "lanes_active": val(r, "smsp__thread_inst_executed_per_inst_executed.ratio"),
        "warps_active_pct": val(r, "sm__warps_active.avg.pct_of_peak_sustained_active"),
        "registers": val(r, "launch__registers_per_thread"),
        "formula": f"smsp__inst_executed.sum / ({sms} SMs x 4 schedulers x sm__cycles_elapsed.max)",
        "source": f"profiles/ (ncu --set full --clock-control none; key {key})",
    },
}
print(json.dumps({key: entry}))
