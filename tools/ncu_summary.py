#!/usr/bin/env python
"""Turn an `ncu --set full` report into the summaries that are committed under profiles/.

    python tools/ncu_summary.py gpurun_out/r01d_prof.ncu-rep profiles/r01d
    python tools/ncu_summary.py gpurun_out/r02n_ncu_store.csv profiles/r02n_store [command line that was profiled]
        (a .csv argument is the output of `ncu -i report --page raw --csv`, made on the GPU box)

writes  profiles/r01d_ncu_full.csv   one row per captured launch, the metrics that matter here
        profiles/ncu_traffic.json    DRAM bytes per full pass of the store-mode Fitch kernel
                                     (bench.py's roofline.traffic reads this file)
Runs on the CPU box (ncu -i reads the report; no GPU needed).
"""
from __future__ import annotations

import csv
import io
import json
import os
import subprocess
import sys

KEEP = [
    "ID", "Kernel Name", "Grid Size", "Block Size",
    "gpu__time_duration.sum",
    "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__bytes_read.sum.per_second", "dram__bytes_write.sum.per_second",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "launch__shared_mem_per_block_dynamic",
    "sm__cycles_elapsed.avg.per_second",
    "sm__inst_issued.avg.pct_of_peak_sustained_active", "smsp__inst_issued.sum",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "smsp__average_warp_latency_per_inst_issued.ratio", "smsp__warps_eligible.avg.per_cycle_active",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
]

SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def main() -> None:
    rep, prefix = sys.argv[1], sys.argv[2]
    command = sys.argv[3] if len(sys.argv) > 3 else None
    if rep.endswith(".csv"):
        raw = open(rep).read()
    else:
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    cols = [hdr.index(k) for k in KEEP if k in hdr]
    out = prefix + "_ncu_full.csv"
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow([hdr[i] for i in cols])
        w.writerow([units[i] for i in cols])
        for r in data:
            w.writerow([r[i] for i in cols])
    print("wrote", out, len(data), "launches")

    def val(r, name):
        i = hdr.index(name)
        return float(r[i]) * SCALE.get(units[i], 1.0)

    tpath = os.path.join(os.path.dirname(out), "ncu_traffic.json")
    tj = json.load(open(tpath)) if os.path.exists(tpath) else {}
    rel = os.path.relpath(out, os.path.dirname(os.path.dirname(os.path.abspath(out))))
    name = lambda r: r[hdr.index("Kernel Name")]
    stream = [r for r in data if "fitch_stream_kernel" in name(r)]
    if stream:
        rd = sum(val(r, "dram__bytes_read.sum") for r in stream) / len(stream)
        wr = sum(val(r, "dram__bytes_write.sum") for r in stream) / len(stream)
        us = sum(float(r[hdr.index("gpu__time_duration.sum")]) for r in stream) / len(stream)
        tj.update({"kernel": "fitch_stream_kernel", "grid": stream[0][hdr.index("Grid Size")], "launches_averaged": len(stream),
                   "dram_bytes_read_per_pass": rd, "dram_bytes_write_per_pass": wr, "dram_bytes_per_pass": rd + wr,
                   "duration_us_under_ncu": us, "source": rel, "command": command})
    score = [r for r in data if name(r).startswith("fitch_score2_kernel")] or \
            [r for r in data if name(r).startswith("fitch_score_kernel")]
    if score:
        rd = sum(val(r, "dram__bytes_read.sum") for r in score) / len(score)
        wr = sum(val(r, "dram__bytes_write.sum") for r in score) / len(score)
        pick = lambda k: sum(float(r[hdr.index(k)]) for r in score) / len(score)
        tj.update({"score_kernel": name(score[0]).split("(")[0] + " (512 NNI candidates of a 2000 x 1M tree)",
                   "score_dram_bytes_per_batch": rd + wr, "score_duration_us_under_ncu": pick("gpu__time_duration.sum"),
                   "score_l2_hit_pct": pick("lts__t_sector_hit_rate.pct"),
                   "score_alu_pipe_pct": pick("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active"),
                   "score_issue_pct": pick("sm__inst_issued.avg.pct_of_peak_sustained_active"),
                   "score_source": rel, "score_command": command})
    nni = [r for r in data if "nni_neighbourhood_kernel" in name(r)]
    if nni:
        pick = lambda k: sum(float(r[hdr.index(k)]) for r in nni) / len(nni)
        tj.update({"nni_dram_bytes_per_call": sum(val(r, "dram__bytes_read.sum") + val(r, "dram__bytes_write.sum") for r in nni) / len(nni),
                   "nni_duration_us_under_ncu": pick("gpu__time_duration.sum"),
                   "nni_alu_pipe_pct": pick("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active"),
                   "nni_issue_pct": pick("sm__inst_issued.avg.pct_of_peak_sustained_active"), "nni_source": rel})
    if stream or score or nni:
        json.dump(tj, open(tpath, "w"), indent=1)
        print("wrote", tpath, tj)
    store = [r for r in data if "fitch_exec_kernel<1>" in r[hdr.index("Kernel Name")] or
             "fitch_exec_kernelILb1" in r[hdr.index("Kernel Name")]]
    if store and not stream:
        # full passes of the bench's default tree have the same grid; take the most common grid
        grids = [r[hdr.index("Grid Size")] for r in store]
        g = max(set(grids), key=grids.count)
        sel = [r for r in store if r[hdr.index("Grid Size")] == g]
        rd = sum(val(r, "dram__bytes_read.sum") for r in sel) / len(sel)
        wr = sum(val(r, "dram__bytes_write.sum") for r in sel) / len(sel)
        us = sum(float(r[hdr.index("gpu__time_duration.sum")]) for r in sel) / len(sel)
        tj = {
            "kernel": "fitch_exec_kernel<true>", "grid": g, "launches_averaged": len(sel),
            "dram_bytes_read_per_pass": rd, "dram_bytes_write_per_pass": wr, "dram_bytes_per_pass": rd + wr,
            "duration_us_under_ncu": us, "source": os.path.relpath(out, os.path.dirname(os.path.dirname(os.path.abspath(out)))),
            "command": "ncu --set full --clock-control none --import-source on -k regex:fitch_exec "
                       "python bench.py --steps 5 --warmup 3 --no-cpu-baseline --candidates 64",
        }
        tpath = os.path.join(os.path.dirname(out), "ncu_traffic.json")
        json.dump(tj, open(tpath, "w"), indent=1)
        print("wrote", tpath, tj)


if __name__ == "__main__":
    main()
