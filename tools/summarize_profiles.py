#!/usr/bin/env python
"""Turns the ncu artefacts a round brought back in gpurun_out/ into the committed summaries under profiles/.

  python tools/summarize_profiles.py r02
    gpurun_out/r02_launches.csv     (ncu --metrics gpu__time_duration.sum ... --csv of bench.py)  -> profiles/r02_launch_list_summary.csv
    gpurun_out/r02_kernels.ncu-rep  (ncu --set full of tools/profile_driver.py)                    -> profiles/r02_kernels_ncu_summary.json
                                                                                                   -> profiles/r02_expand_steady_ncu_summary.json
"""
import csv
import json
import os
import subprocess
import sys
from collections import defaultdict

rnd = sys.argv[1] if len(sys.argv) > 1 else "r02"


def short(name):
    return name.split("(")[0].replace("void ", "").replace("gs::", "").replace("(anonymous namespace)::", "")


launch_csv = f"gpurun_out/{rnd}_launches.csv"
if os.path.exists(launch_csv):
    rows = list(csv.reader(open(launch_csv)))
    for i, r in enumerate(rows):
        if "Kernel Name" in r:
            hdr, start = r, i
            break
    ki, mi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    d = defaultdict(list)
    for r in rows[start + 1:]:
        if len(r) > mi:
            try:
                d[short(r[ki])].append(float(r[mi].replace(",", "")))
            except ValueError:
                pass
    lines = ["kernel,launches,mean_us,min_us,max_us,total_us"]
    for k, v in sorted(d.items(), key=lambda kv: -sum(kv[1])):
        lines.append(f"{k},{len(v)},{sum(v)/len(v)/1e3:.2f},{min(v)/1e3:.2f},{max(v)/1e3:.2f},{sum(v)/1e3:.1f}")
    open(f"profiles/{rnd}_launch_list_summary.csv", "w").write("\n".join(lines) + "\n")
    print("\n".join(lines))

rep = f"gpurun_out/{rnd}_kernels.ncu-rep"
if os.path.exists(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rr = list(csv.reader(raw.splitlines()))
    h, u = rr[0], rr[1]
    want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sectors_op_write.sum",
            "lts__t_sectors_op_red.sum", "lts__t_sectors_op_atom.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "smsp__inst_executed.sum", "sm__cycles_elapsed.max",
            "sm__cycles_active.avg", "sm__inst_executed.avg.per_cycle_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
            "lts__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers", "smsp__inst_executed_op_global_st.sum",
            "smsp__inst_executed_op_global_red.sum", "smsp__inst_executed_op_global_atom.sum"]
    stall = [c for c in h if c.startswith("smsp__average_warps_issue_stalled_") and c.endswith("_per_issue_active.ratio")]

    def num(v, unit, name):
        f = float(v.replace(",", ""))
        if "bytes" in name:
            f *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)
        if name == "gpu__time_duration.sum":
            f *= {"ns": 1e-3, "us": 1, "ms": 1e3, "usecond": 1, "nsecond": 1e-3, "msecond": 1e3}.get(unit, 1)
        return f

    per = defaultdict(list)
    for r in rr[2:]:
        m = {}
        for w in want + stall:
            if w in h:
                i = h.index(w)
                try:
                    m[w] = num(r[i], u[i], w)
                except ValueError:
                    pass
        m["dram_bytes"] = m.get("dram__bytes_read.sum", 0) + m.get("dram__bytes_write.sum", 0)
        per[short(r[h.index("Kernel Name")])].append(m)
    out = {"round": rnd, "command": "ncu --set full --clock-control none --cache-control none --import-source on python tools/profile_driver.py",
           "note": "cache control off: the L2 is not flushed between launches, so a launch sees the state the launches before it left; "
                   "times under ncu are serialised (no overlap between launches): compare shares and counters, not absolutes",
           "kernels": {}}
    for k, ls in per.items():
        last = ls[-1]
        top = sorted(((c.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""), last[c]) for c in stall if c in last),
                     key=lambda kv: -kv[1])[:5]
        out["kernels"][k] = {"launches_captured": len(ls), "time_us": [round(m["gpu__time_duration.sum"], 2) for m in ls],
                             "dram_bytes": [int(m["dram_bytes"]) for m in ls], "last_launch": {w: last.get(w) for w in want if w in last},
                             "top_stalls_last_launch(warps per issue)": top}
    json.dump(out, open(f"profiles/{rnd}_kernels_ncu_summary.json", "w"), indent=1)
    exp = [k for k in per if k.startswith("k_expand")]
    if exp:
        ls = per[exp[0]]
        steady = ls[8:] if len(ls) > 8 else ls[-1:]
        json.dump({"round": rnd, "kernel": exp[0], "launches": "9th and later of twelve pipelined batches into four rotating outputs, cache control off",
                   "dram_bytes_per_launch": sum(m["dram_bytes"] for m in steady) / len(steady),
                   "dram_bytes_write_per_launch": sum(m.get("dram__bytes_write.sum", 0) for m in steady) / len(steady),
                   "time_us": [round(m["gpu__time_duration.sum"], 2) for m in steady]},
                  open(f"profiles/{rnd}_expand_steady_ncu_summary.json", "w"), indent=1)
    for k, v in out["kernels"].items():
        print(k, v["time_us"], v["dram_bytes"][-1], v["top_stalls_last_launch(warps per issue)"][:3])
