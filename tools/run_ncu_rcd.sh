#!/bin/bash
# one ncu pass over the relu_concat_dropout kernels (after the plain run has exited 0)
set -e
mkdir -p gpurun_out
python tools/ab_relu_concat_dropout.py 515000 > gpurun_out/rcd_ab.log 2>&1
ncu --clock-control none -k regex:k_rcd_${1:-backward} --launch-count 2 --launch-skip 3 \
  --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__grid_size,l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum,l1tex__t_requests_pipe_lsu_mem_global_op_st.sum,l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum,l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum,lts__t_sectors_srcunit_tex_op_write.sum,lts__t_sectors_srcunit_tex_op_read.sum,lts__t_sectors_op_read.sum,lts__t_sectors_op_write.sum,sm__inst_executed_pipe_lsu.sum,smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct,smsp__warp_issue_stalled_lg_throttle_per_warp_active.pct,smsp__warp_issue_stalled_drain_per_warp_active.pct \
  --csv --log-file gpurun_out/rcd_ncu.csv python tools/ab_relu_concat_dropout.py 515000 > gpurun_out/rcd_ncu.log 2>&1
