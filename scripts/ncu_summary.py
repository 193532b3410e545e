#!/usr/bin/env python
"""Prints the metrics the judge asks for from an .ncu-rep (read here, no GPU):  python scripts/ncu_summary.py REPORT [out.md]"""
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__occupancy_limit_registers",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__t_bytes.sum", "l1tex__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum",
    "l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "smsp__sass_average_branch_targets_threads_uniform.pct", "sm__sass_branch_targets_threads_divergent.sum", "sm__sass_branch_targets.sum",
    "smsp__warps_eligible.avg.per_cycle_active", "smsp__warps_active.avg.per_cycle_active",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_imc_miss_per_issue_active.ratio",
]


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    out = []
    for r in rows[2:]:
        get = lambda k: r[hdr.index(k)] if k in hdr else None
        out.append(f"## {get('Kernel Name')[:90]}  grid {get('Grid Size')} block {get('Block Size')}")
        out.append("| metric | value | unit |\n|---|---|---|")
        for k in KEYS:
            if k in hdr:
                out.append(f"| {k} | {get(k)} | {units[hdr.index(k)]} |")
        s, q = get("l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum"), get("l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum")
        if s and q:
            out.append(f"| sectors per request (global loads) | {float(s.replace(',', '')) / float(q.replace(',', '')):.2f} | |")
        t = get("gpu__time_duration.sum")
        if t:
            tn = float(t.replace(",", "")) * {"ns": 1e-9, "us": 1e-6, "usecond": 1e-6, "ms": 1e-3, "msecond": 1e-3, "nsecond": 1e-9}.get(units[hdr.index("gpu__time_duration.sum")], 1e-9)
            for nm in ("lts__t_bytes.sum", "l1tex__t_bytes.sum"):
                v = get(nm)
                if v:
                    mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(units[hdr.index(nm)], 1)
                    out.append(f"| {nm} rate | {float(v.replace(',', '')) * mult / tn / 1e9:.1f} | GB/s |")
            dr, dw = get("dram__bytes_read.sum"), get("dram__bytes_write.sum")
            if dr and dw:
                m1 = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(units[hdr.index("dram__bytes_read.sum")], 1)
                m2 = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(units[hdr.index("dram__bytes_write.sum")], 1)
                tot = float(dr.replace(",", "")) * m1 + float(dw.replace(",", "")) * m2
                out.append(f"| dram bytes read+write | {tot:.0f} | byte |")
                out.append(f"| dram rate | {tot / tn / 1e9:.1f} | GB/s |")
        out.append("")
    text = "\n".join(out)
    print(text)
    if len(sys.argv) > 2:
        open(sys.argv[2], "w").write(text + "\n")


if __name__ == "__main__":
    main()
