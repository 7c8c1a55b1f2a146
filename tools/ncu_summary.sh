#!/bin/bash
# usage: tools/ncu_summary.sh <report.ncu-rep>  -> key raw metrics per captured launch (CSV transposed to text)
ncu -i "$1" --page raw --csv 2>/dev/null | python3 -c '
import csv,sys
rows=list(csv.reader(sys.stdin))
hdr=rows[0]; units=rows[1]; data=rows[2:]
keys=["Kernel Name","Grid Size","Block Size","gpu__time_duration.sum","launch__registers_per_thread","launch__occupancy_limit_registers","launch__occupancy_limit_shared_mem","sm__warps_active.avg.pct_of_peak_sustained_active","sm__throughput.avg.pct_of_peak_sustained_elapsed","gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed","dram__bytes_read.sum","dram__bytes_write.sum","lts__t_sector_hit_rate.pct","l1tex__t_sector_hit_rate.pct","lts__throughput.avg.pct_of_peak_sustained_elapsed","l1tex__throughput.avg.pct_of_peak_sustained_elapsed","sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active","sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active","sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active","sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active","sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active","sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active","smsp__issue_active.avg.pct_of_peak_sustained_active","smsp__inst_executed.sum","smsp__thread_inst_executed_per_inst_executed.ratio","smsp__thread_inst_executed_per_inst_executed.pct","sm__average_thread_inst_executed_pred_on_per_inst_executed_realtime.ratio","smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct","smsp__warp_issue_stalled_short_scoreboard_per_warp_active.pct","smsp__warp_issue_stalled_wait_per_warp_active.pct","smsp__warp_issue_stalled_barrier_per_warp_active.pct","smsp__warp_issue_stalled_math_pipe_throttle_per_warp_active.pct","smsp__warp_issue_stalled_branch_resolving_per_warp_active.pct","smsp__warp_issue_stalled_no_instruction_per_warp_active.pct","smsp__warp_issue_stalled_not_selected_per_warp_active.pct","smsp__warp_issue_stalled_lg_throttle_per_warp_active.pct","smsp__warp_issue_stalled_mio_throttle_per_warp_active.pct","smsp__warp_issue_stalled_dispatch_stall_per_warp_active.pct","smsp__warp_issue_stalled_membar_per_warp_active.pct","sm__cycles_active.avg","smsp__cycles_active.avg","local_load","smsp__inst_executed_op_local_ld.sum","smsp__inst_executed_op_local_st.sum"]
for k in keys:
    idx=[i for i,h in enumerate(hdr) if h==k]
    if not idx: continue
    i=idx[0]
    print("%-78s %-10s %s"%(k,units[i]," | ".join(r[i][:40] for r in data)))
'
