#!/usr/bin/env python3
"""Summarise ncu output into small tracked files under profiles/ (run here, no GPU needed):
    python tools/ncu_summary.py launches gpurun_out/X_launches.csv profiles/NAME.json
    python tools/ncu_summary.py full gpurun_out/X.ncu-rep profiles/NAME.json
"""
import collections
import csv
import io
import json
import os
import subprocess
import sys


def launches(src, dst):
    rows = [r for r in csv.reader(open(src)) if len(r) > 5]
    hdr = rows[0]
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = collections.OrderedDict()
    seq = []
    for r in rows[1:]:
        name = r[ki].split("(")[0].replace("void ", "").replace("gprt::<unnamed>::", "").strip()
        ns = float(r[vi].replace(",", ""))
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += ns
        seq.append((name, ns))
    total = sum(v[1] for v in agg.values())
    out = {"source": src, "note": "ncu --metrics gpu__time_duration.sum --clock-control none: cold-cache, serialised launches; compare SHARES",
           "total_us": round(total / 1e3, 1),
           "kernels": [{"kernel": k, "launches": v[0], "total_us": round(v[1] / 1e3, 1), "avg_us": round(v[1] / v[0] / 1e3, 2),
                        "share": round(v[1] / total, 4)} for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])]}
    # one frame of the render loop: from a primary traverse to the final fold
    frame = []
    for name, ns in seq:
        if name.startswith(("k_walk<1", "k_walk_deep<1", "k_traverse_brute<1")):  # the primary traversal opens a frame
            frame = []
        frame.append((name, round(ns / 1e3, 1)))
        if name.startswith("k_final_color") and len(frame) >= 13:
            break
    out["one_frame_sequence_us"] = frame
    json.dump(out, open(dst, "w"), indent=1)
    print(json.dumps(out["kernels"][:8], indent=1))


WANT = ["gpu__time_duration.sum", "lts__t_sectors_op_read.sum", "lts__t_sectors_op_write.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__issue_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed.avg.per_cycle_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__thread_inst_executed_per_inst_executed.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "derived__smsp__sass_thread_inst_executed_op_ffma_pred_on_x2", "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum",
        "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum", "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum",
        "smsp__inst_executed.sum", "smsp__thread_inst_executed.sum", "lts__t_bytes.sum", "l1tex__t_bytes.sum",
        "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "smsp__average_warp_latency_per_inst_issued.ratio"]


def full(src, dst):
    txt = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr = rows[0]
    units = dict(zip(hdr, rows[1]))
    out = {"source": src, "made_by": " ".join(["python", "tools/ncu_summary.py"] + sys.argv[1:]) + "  (the .ncu-rep is scratch under gpurun_out/ and is not tracked: re-capture it with the "
                                     "command in the `capture` field on a B200, then re-run this line)",
           "capture": os.environ.get("NCU_CAPTURE_CMD", ""),
           "note": "ncu --set full --clock-control none (replayed, cold cache); one entry per captured launch",
           "units": {k: units.get(k, "") for k in WANT if k in units}, "launches": []}
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        e = {"kernel": d.get("Kernel Name", "")[:80]}
        for k in WANT:
            if k in d and d[k] != "":
                try:
                    e[k] = float(d[k].replace(",", ""))
                except ValueError:
                    e[k] = d[k]
        if "smsp__thread_inst_executed_per_inst_executed.ratio" in e:
            e["lane_efficiency"] = round(e["smsp__thread_inst_executed_per_inst_executed.ratio"] / 32, 3)
        if "dram__bytes_read.sum" in e:
            scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
            e["dram_bytes_total"] = e["dram__bytes_read.sum"] * scale.get(units.get("dram__bytes_read.sum", "byte"), 1.0) + \
                e.get("dram__bytes_write.sum", 0.0) * scale.get(units.get("dram__bytes_write.sum", "byte"), 1.0)
        out["launches"].append(e)
    json.dump(out, open(dst, "w"), indent=1)
    for e in out["launches"]:
        print({k: e[k] for k in ("gpu__time_duration.sum", "dram_bytes_total", "lane_efficiency", "sm__issue_active.avg.pct_of_peak_sustained_elapsed",
                                 "launch__registers_per_thread") if k in e})


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3])
