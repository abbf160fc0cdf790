#!/usr/bin/env python
"""Turn an Nsight Compute report (.ncu-rep, captured with --set full) into the short text
summary committed under profiles/:  python tools/ncu_summary.py gpurun_out/x.ncu-rep > profiles/x.txt
Also summarises a launch list CSV (--metrics gpu__time_duration.sum):  ... launches.csv"""
import collections
import csv
import subprocess
import sys

WANT = [
    "gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum", "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum",
    "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum",
    "smsp__sass_thread_inst_executed_op_dfma_pred_on.sum",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_bytes.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
]


def report(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for row in rows[2:]:
        rec = dict(zip(hdr, row))
        print("kernel:", rec.get("Kernel Name"))
        for k in WANT:
            if k in rec:
                print("  %-82s %-14s %s" % (k, units[hdr.index(k)], rec[k]))
        t = float(rec["gpu__time_duration.sum"].replace(",", ""))
        tu = units[hdr.index("gpu__time_duration.sum")]
        sec = t * {"ns": 1e-9, "us": 1e-6, "usecond": 1e-6, "ms": 1e-3, "msecond": 1e-3, "nsecond": 1e-9, "second": 1}.get(tu, 1e-3)
        fl = 0.0
        for op, w in (("fadd", 1), ("fmul", 1), ("ffma", 2)):
            key = "smsp__sass_thread_inst_executed_op_%s_pred_on.sum" % op
            if key in rec:
                fl += w * float(rec[key].replace(",", ""))
        if fl:
            print("  executed FP32 flop rate (fadd + fmul + 2 ffma) / duration: %.2f TFLOP/s" % (fl / sec / 1e12))
        rd = float(rec.get("dram__bytes_read.sum", "0").replace(",", ""))
        wr = float(rec.get("dram__bytes_write.sum", "0").replace(",", ""))
        print("  dram traffic per launch (read + write, units as above): %g + %g" % (rd, wr))


def launches(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 10]
    hdr = rows[0]
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = collections.defaultdict(list)
    for r in rows[1:]:
        agg[r[ki]].append(float(r[vi].replace(",", "")))
    tot = sum(sum(v) for v in agg.values())
    print("launch list %s: %d launches (cold-cache, serialised: compare shares, not absolutes)" % (path, len(rows) - 1))
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        print("  %-70s n=%3d mean=%10.1f us share=%5.1f%%" % (k[:70], len(v), sum(v) / len(v) / 1e3, 100 * sum(v) / tot))


if __name__ == "__main__":
    for p in sys.argv[1:]:
        (launches if p.endswith(".csv") else report)(p)
