#!/bin/bash
# GPU-box helper: a plain run first (exit 0 required), then targeted ncu metrics for kernels matching $1
# (skip $2 launches, capture $3), CSV to gpurun_out/metrics.csv.
mkdir -p gpurun_out
M=gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum
M=$M,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,smsp__issue_active.avg.pct_of_peak_sustained_active
M=$M,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,lts__throughput.avg.pct_of_peak_sustained_elapsed
M=$M,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio
M=$M,smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio
M=$M,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio
M=$M,smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio
M=$M,smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio
M=$M,smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio
M=$M,smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio
M=$M,l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum,l1tex__data_pipe_lsu_wavefronts.sum
CMD="python bench.py --steps 8 --warmup 8 --no-cpu --no-e2e ${4:-}"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics $M --clock-control none -k regex:$1 -s $2 -c $3 --csv \
  --log-file gpurun_out/metrics.csv $CMD > gpurun_out/ncu_metrics.log 2>&1
tail -1 gpurun_out/ncu_metrics.log | cut -c1-120
