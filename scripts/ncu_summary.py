"""Compact per-launch table from `ncu -i X.ncu-rep --page raw --csv` (run on the GPU box so only the summary travels)."""
import csv
import sys

KEYS = [
    ("gpu__time_duration.sum", "time"),
    ("dram__bytes_read.sum", "dram_rd"),
    ("dram__bytes_write.sum", "dram_wr"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
    ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1%"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
    ("smsp__issue_active.avg.pct", "issue%"),
    ("smsp__inst_executed.sum", "winst"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__occupancy_limit_registers", "lim_reg"),
    ("launch__occupancy_limit_shared_mem", "lim_smem"),
    ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "fma%"),
    ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "fp64%"),
    ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu%"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor%"),
    ("sm__inst_executed_pipe_tc.avg.pct_of_peak_sustained_active", "tc%"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "bank_conf"),
    ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "st_long"),
    ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "st_short"),
    ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "st_barrier"),
    ("smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio", "st_mio"),
    ("smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "st_lg"),
    ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "st_wait"),
    ("smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "st_math"),
    ("smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "st_notsel"),
    ("smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio", "st_noinst"),
]
csv.field_size_limit(1 << 30)
rows = list(csv.reader(open(sys.argv[1], errors="replace")))
rows = [r for r in rows if len(r) > 10]
hdr, units = rows[0], rows[1]
col = {h: i for i, h in enumerate(hdr)}
name_i = col.get("Kernel Name")
print("| kernel | " + " | ".join(k for _, k in KEYS) + " |")
print("|---|" + "---|" * len(KEYS))
for r in rows[2:]:
    name = r[name_i].split("(")[0].replace("void ", "").replace("<unnamed>::", "")[:48]
    vals = []
    for h, _ in KEYS:
        i = col.get(h)
        if i is None:
            vals.append("-")
            continue
        v, u = r[i], units[i]
        try:
            f = float(v.replace(",", ""))
            v = f"{f:.4g}"
        except ValueError:
            pass
        vals.append(f"{v} {u}".strip() if u in ("ms", "us", "ns", "Gbyte", "Mbyte", "Kbyte", "byte") else v)
    print(f"| {name} | " + " | ".join(vals) + " |")
