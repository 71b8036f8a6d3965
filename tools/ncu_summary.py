"""Summarise an `ncu --page raw --csv` export: one line per launch with the metrics the roofline block needs."""
import csv, sys, json
WANT = {"gpu__time_duration.sum": "ns", "smsp__inst_executed.sum": "warp_inst", "dram__bytes_read.sum": "dram_rd", "dram__bytes_write.sum": "dram_wr",
        "smsp__issue_active.avg.per_cycle_active": "issue_active", "smsp__warps_active.avg.per_cycle_active": "warps_active_per_sched",
        "launch__registers_per_thread": "regs", "launch__grid_size": "grid", "launch__block_size": "block", "launch__shared_mem_per_block_dynamic": "smem_dyn",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_pct", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram_pct",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio": "stall_long_sb",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio": "stall_short_sb", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio": "stall_wait",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio": "stall_barrier", "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio": "stall_mio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio": "stall_lg", "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio": "stall_math",
        "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio": "stall_noinst", "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio": "stall_branch",
        "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio": "stall_dispatch", "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio": "stall_notsel",
        "sm__warps_active.avg.pct_of_peak_sustained_active": "occupancy_pct"}


def rows(path):
    r = list(csv.reader(open(path)))
    hdr = next(i for i, x in enumerate(r) if x and x[0] == "ID")
    names, units = r[hdr], r[hdr + 1]
    out = []
    for x in r[hdr + 2:]:
        if len(x) != len(names):
            continue
        d = {"kernel": x[names.index("Kernel Name")].split("(")[0]}
        for n, short in WANT.items():
            if n in names:
                v = x[names.index(n)].replace(",", "")
                try:
                    d[short] = float(v)
                except ValueError:
                    d[short] = v
                u = units[names.index(n)].lower()
                if short == "ns":
                    d[short] *= {"us": 1e3, "ms": 1e6, "s": 1e9}.get(u, 1)
                if short in ("dram_rd", "dram_wr"):
                    d[short] *= {"kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
        out.append(d)
    return out


if __name__ == "__main__":
    for d in rows(sys.argv[1]):
        print(json.dumps(d))
