#!/usr/bin/env python
"""Print the handful of ncu metrics the profiles/ notes quote from a .ncu-rep (reads it with `ncu -i ... --page raw --csv`)."""
import csv, subprocess, sys
WANT = """gpu__time_duration.sum launch__registers_per_thread launch__grid_size launch__block_size launch__shared_mem_per_block_static
launch__shared_mem_per_block_dynamic launch__occupancy_limit_registers launch__occupancy_limit_shared_mem sm__warps_active.avg.pct_of_peak_sustained_active
smsp__thread_inst_executed_per_inst_executed.ratio smsp__inst_executed.sum dram__bytes_read.sum dram__bytes_write.sum
dram__throughput.avg.pct_of_peak_sustained_elapsed lts__t_sectors.sum lts__t_sector_hit_rate.pct l1tex__t_sector_hit_rate.pct
lts__throughput.avg.pct_of_peak_sustained_elapsed l1tex__throughput.avg.pct_of_peak_sustained_elapsed sm__throughput.avg.pct_of_peak_sustained_elapsed
smsp__issue_active.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active
smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio
smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio smsp__average_warps_issue_stalled_wait_per_issue_active.ratio
smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio
smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio
smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio
smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio smsp__average_warps_issue_stalled_membar_per_issue_active.ratio
smsp__average_warps_issue_stalled_imc_miss_per_issue_active.ratio smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio
l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum smsp__inst_executed_op_shared_ld.sum smsp__inst_executed_op_shared_st.sum
smsp__inst_executed_op_global_ld.sum smsp__inst_executed_op_global_st.sum smsp__inst_executed_op_local_ld.sum smsp__inst_executed_op_local_st.sum""".split()
for rep in sys.argv[1:]:
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(out.splitlines()))
    h, u = r[0], r[1]
    for row in r[2:]:
        print("==", rep, row[h.index("Kernel Name")][:60])
        for w in WANT:
            if w in h:
                i = h.index(w); print(f"{w:82s} {u[i]:>10s} {row[i]}")
