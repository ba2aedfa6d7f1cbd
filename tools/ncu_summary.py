#!/usr/bin/env python
"""Summarises an `ncu --page raw --csv` export (and optionally a `--page source --csv` one) into a small text file for profiles/.
    python tools/ncu_summary.py raw.csv [source.csv] > profiles/<name>.txt"""
import csv
import sys

WANT = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__occupancy_limit_registers",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__sectors_read.sum",
    "dram__sectors_read.sum.per_second", "dram__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "lts__t_requests_srcunit_tex_op_read.sum", "lts__t_sectors_srcunit_tex_op_read.sum",
    "lts__t_sectors_srcunit_tex_op_read_lookup_miss.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__t_sector_hit_rate.pct", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
    "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum", "smsp__issue_active.avg.per_cycle_active",
    "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__warps_active.avg.per_cycle_active", "smsp__warps_eligible.avg.per_cycle_active",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
]


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    for r in rows[2:]:
        print("=" * 100)
        print("kernel:", r[col["Kernel Name"]])
        for w in WANT:
            if w in col and r[col[w]] != "":
                print(f"  {w:90s} {r[col[w]]:>22s} {units[col[w]]}")
        stalls = []
        for h, i in col.items():
            if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio"):
                try:
                    stalls.append((float(r[i]), h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]))
                except ValueError:
                    pass
        tot = sum(v for v, _ in stalls) or 1.0
        print("  stall reasons (warps per issue slot, share):", ", ".join(f"{n} {v:.2f} ({100 * v / tot:.0f}%)" for v, n in sorted(stalls, reverse=True)[:6]))
    if len(sys.argv) > 2:
        rows = list(csv.reader(open(sys.argv[2])))
        hdr, kern, best, seen = None, "?", {}, set()
        for r in rows:
            if len(r) >= 2 and r[0] == "Kernel Name":
                kern = r[1].split("(")[0][-40:]
                continue
            if "Source" in r and "# Samples" in r:
                hdr = r
                continue
            if hdr is None or len(r) != len(hdr):
                continue
            c = {h: i for i, h in enumerate(hdr)}
            if (kern, r[c["Address"]]) in seen:
                continue
            seen.add((kern, r[c["Address"]]))
            try:
                best.setdefault(kern, []).append((float(r[c["# Samples"]].replace(",", "")), r[c["Source"]].strip()[:110],
                                                  r[c["Instructions Executed"]], r[c["Avg. Threads Executed"]]))
            except (ValueError, KeyError):
                pass
        for kern, lines in best.items():
            print("=" * 100)
            print(f"hottest SASS instructions by warp-stall samples: {kern}   (share, instruction, executed, avg active threads)")
            tot = sum(v for v, *_ in lines) or 1.0
            for v, src, ex, th in sorted(lines, reverse=True)[:18]:
                print(f"  {100 * v / tot:5.1f}%  {src:110s} {ex:>12s} {th:>4s}")


if __name__ == "__main__":
    main()
