#!/usr/bin/env python
"""Summarise an ncu report (run HERE, no GPU needed): selected raw metrics per launch + the hot loop of the source page.
Usage: summarize_ncu.py REPORT.ncu-rep OUT_PREFIX   ->  OUT_PREFIX.metrics.txt, OUT_PREFIX.hotloop.txt"""
import csv
import io
import subprocess
import sys

KEEP = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__cycles_active.avg", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__issue_active.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__sass_thread_inst_executed_op_dfma_pred_on.sum", "smsp__sass_thread_inst_executed_op_dmul_pred_on.sum",
        "smsp__sass_thread_inst_executed_op_dadd_pred_on.sum", "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum",
        "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum", "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum",
        "smsp__sass_thread_inst_executed_op_dfma_pred_on.sum.per_cycle_elapsed", "smsp__sass_thread_inst_executed_op_dmul_pred_on.sum.per_cycle_elapsed",
        "smsp__sass_thread_inst_executed_op_dadd_pred_on.sum.per_cycle_elapsed",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sectors_op_read.sum", "lts__t_sectors_op_write.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio"]


def page(rep, name):
    out = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep, prefix = sys.argv[1], sys.argv[2]
    rows = page(rep, "raw")
    hdr, units = rows[0], rows[1]
    with open(prefix + ".metrics.txt", "w") as f:
        f.write(f"# ncu --set full --clock-control none, report {rep.split('/')[-1]}; per launch\n")
        for r in rows[2:]:
            d = dict(zip(hdr, r))
            u = dict(zip(hdr, units))
            f.write(f"\n## launch {d.get('ID')}: {d.get('Kernel Name')}  grid {d.get('Grid Size')} block {d.get('Block Size')}\n")
            for k in KEEP:
                if k in d:
                    f.write(f"{k:95s} {d[k]:>18s} {u.get(k, '')}\n")
    rows = page(rep, "source")
    hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
    if not hi:
        return
    h = rows[hi[0]]
    end = hi[1] - 1 if len(hi) > 1 else len(rows)
    col = {k: i for i, k in enumerate(h)}
    body = [r for r in rows[hi[0] + 1:end] if len(r) > 10]
    mx = max(int(r[col["Instructions Executed"]] or 0) for r in body)
    tot = sum(int(r[col["# Samples"]] or 0) for r in body)
    with open(prefix + ".hotloop.txt", "w") as f:
        f.write(f"# {rows[0][1] if rows and len(rows[0]) > 1 else ''}\n# SASS of the hot loop (instructions executed >= 90% of the max), warp-stall samples per instruction; total samples {tot}\n")
        f.write(f"{'addr':>6s} {'SASS':60s} {'samples':>8s} {'math':>7s} {'wait':>6s} {'notsel':>7s} {'disp':>6s} {'ssb':>6s}\n")
        cnt = {}
        for r in body:
            if int(r[col["Instructions Executed"]] or 0) >= 0.9 * mx:
                src = r[col["Source"]]
                op = src.split()[1 if src.startswith("@") else 0].split(".")[0]
                cnt[op] = cnt.get(op, 0) + 1
                f.write(f"{r[col['Address']][-5:]:>6s} {src[:60]:60s} {r[col['# Samples']]:>8s} {r[col['stall_math']]:>7s} {r[col['stall_wait']]:>6s} "
                        f"{r[col['stall_not_selected']]:>7s} {r[col['stall_dispatch']]:>6s} {r[col['stall_short_sb']]:>6s}\n")
        f.write("\n# opcode histogram of the hot loop: " + ", ".join(f"{k} {v}" for k, v in sorted(cnt.items(), key=lambda kv: -kv[1])) + "\n")


if __name__ == "__main__":
    main()
