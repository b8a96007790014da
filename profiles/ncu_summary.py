#!/usr/bin/env python
"""Turn an `ncu --set full` report into the committed text summary + profiles/ncu_traffic.json.

    python profiles/ncu_summary.py gpurun_out/frame_full.ncu-rep profiles/r1/ncu_full_summary.md [--traffic]

Reads the report with `ncu -i ... --page raw --csv` (works without a GPU).  --traffic rewrites
profiles/ncu_traffic.json (dram read+write bytes per launch of each frame kernel; bench.py reports it as
roofline.traffic)."""
import csv
import io
import json
import os
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "dram read"),
    ("dram__bytes_write.sum", "dram write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram throughput % of peak"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 throughput % of peak"),
    ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "L1/TEX throughput % of peak"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput % of peak"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate"),
    ("l1tex__t_sector_hit_rate.pct", "L1 hit rate"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy"),
    ("smsp__thread_inst_executed_per_inst_executed.ratio", "warp execution efficiency (threads/inst, of 32)"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy"),
    ("launch__registers_per_thread", "registers/thread"),
    ("launch__shared_mem_per_block_dynamic", "dynamic smem/block"),
    ("launch__occupancy_limit_registers", "occupancy limit: registers (blocks)"),
    ("launch__occupancy_limit_shared_mem", "occupancy limit: smem (blocks)"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("smsp__average_warp_latency_per_inst_issued.ratio", "warp latency per instruction issued"),
    ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall long scoreboard / issue"),
    ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "stall short scoreboard / issue"),
    ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "stall wait / issue"),
    ("smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "stall not selected / issue"),
    ("smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio", "stall branch resolving / issue"),
    ("smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio", "stall no instruction / issue"),
    ("smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "stall math pipe throttle / issue"),
]


def main():
    rep, out = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    name_i = hdr.index("Kernel Name")
    grid_i, block_i = hdr.index("Grid Size"), hdr.index("Block Size")
    lines = [f"# ncu --set full summary of `{os.path.basename(rep)}`", "",
             "One launch per kernel, captured under `ncu --set full --clock-control none --import-source on` (per-launch",
             "times are serialised and cold-cache: compare shares, not absolutes).", ""]
    head = "| metric | " + " | ".join(r[name_i].split("(")[0].replace("void ", "") for r in data) + " |"
    lines += [head, "|---|" + "---|" * len(data)]
    lines.append("| grid x block | " + " | ".join(f"{r[grid_i]} x {r[block_i]}" for r in data) + " |")
    traffic = {}
    for key, label in KEYS:
        if key not in hdr:
            continue
        i = hdr.index(key)
        vals = []
        for r in data:
            try:
                vals.append(f"{float(r[i]):,.2f}")
            except ValueError:
                vals.append(r[i])
        lines.append(f"| {label} [{units[i]}] `{key}` | " + " | ".join(vals) + " |")
    mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    ir, iw = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
    for r in data:
        full = r[name_i].split("(")[0].replace("void ", "")
        k = full.split("<")[0].replace("_kernel", "")
        if k == "trace":        # trace_kernel<WAVE, ...>: 0 = primary wave, 1 = shadow wave (bench.py's kernel names)
            k = "shadow" if full.split("<")[1].split(",")[0].strip() == "1" else "primary"
        traffic[k] = int(float(r[ir]) * mult[units[ir]] + float(r[iw]) * mult[units[iw]])
    lines += ["", "dram read+write bytes per launch: " + json.dumps(traffic)]
    open(out, "w").write("\n".join(lines) + "\n")
    if "--frame-json" in sys.argv:
        # profiles/ncu_frame_kernel.json: what bench.py's roofline.issue / roofline.traffic read (LAST frame_kernel row = warm launch)
        want = sys.argv[sys.argv.index("--frame-json") + 1] if len(sys.argv) > sys.argv.index("--frame-json") + 1 and not sys.argv[sys.argv.index("--frame-json") + 1].startswith("--") else "frame_kernel"
        rows_f = [r for r in data if want in r[name_i].replace("(bool)", "").replace("(int)", "")]
        if rows_f:
            r = rows_f[-1]
            g = lambda key: float(r[hdr.index(key)].replace(",", ""))
            dur_mult = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}
            di = hdr.index("gpu__time_duration.sum")
            frame = {"warp_instructions": int(g("smsp__inst_executed.sum")),
                     "lanes_per_instruction": g("smsp__thread_inst_executed_per_inst_executed.ratio"),
                     "issue_active_pct": g("smsp__issue_active.avg.pct_of_peak_sustained_active"),
                     "dram_bytes": int(float(r[ir]) * mult[units[ir]] + float(r[iw]) * mult[units[iw]]),
                     "duration_us_under_ncu": float(r[di]) * dur_mult.get(units[di], 1.0),
                     "kernel": r[name_i].split("(")[0].replace("void ", ""),
                     "source": os.path.relpath(out, os.path.dirname(os.path.abspath(__file__)))}
            json.dump(frame, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ncu_frame_kernel.json"), "w"), indent=1)
    if "--traffic" in sys.argv:
        p = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ncu_traffic.json")
        traffic["_source"] = os.path.relpath(out, os.path.dirname(p))
        json.dump(traffic, open(p, "w"), indent=1)
    print("\n".join(lines))


if __name__ == "__main__":
    main()
