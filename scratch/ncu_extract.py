#!/usr/bin/env python
"""ncu -i <rep> --page raw --csv -> the handful of numbers the profiles/ summaries and bench.py's tracked kernel file quote."""
import csv, io, json, subprocess, sys
WANT = {
 "gpu__time_duration.sum": "duration_ms", "dram__bytes_read.sum": "dram_read", "dram__bytes_write.sum": "dram_write",
 "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed": "l1_wavefront_pct",
 "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_active_pct",
 "smsp__thread_inst_executed_per_inst_executed.ratio": "lanes_per_inst",
 "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_active_pct",
 "launch__registers_per_thread": "registers", "smsp__inst_executed.sum": "warp_inst",
 "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum": "l1_global_load_sectors", "l1tex__t_sector_hit_rate.pct": "l1_hit_pct",
 "lts__t_sector_hit_rate.pct": "l2_hit_pct", "lts__t_sectors.sum": "l2_sectors",
 "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram_pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed": "l2_pct",
 "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active": "pipe_alu_pct", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active": "pipe_fma_pct",
 "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active": "pipe_xu_pct", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active": "pipe_lsu_pct",
 "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio": "stall_long_scoreboard",
 "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio": "stall_wait",
 "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio": "stall_not_selected",
 "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio": "stall_branch_resolving",
 "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio": "stall_math_pipe_throttle",
 "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio": "stall_no_instruction",
 "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio": "stall_short_scoreboard",
 "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio": "stall_lg_throttle",
 "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio": "stall_mio_throttle",
}
SCALE = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "ms": 1.0, "us": 1e-3, "s": 1e3, "ns": 1e-6}
def extract(rep):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    out = {"kernel": vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else ""}
    for k, name in WANT.items():
        if k in hdr:
            i = hdr.index(k)
            v = float(vals[i].replace(",", ""))
            out[name] = v * SCALE.get(units[i], 1.0)
    return out
if __name__ == "__main__":
    print(json.dumps({r: extract(r) for r in sys.argv[1:]}, indent=1))
