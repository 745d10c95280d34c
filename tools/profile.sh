#!/bin/bash
# tools/profile.sh <what> [tag] — the ncu passes behind profiles/*.  Run on the GPU box through gpurun, e.g.
#   gpurun --timeout 1500 -- 'bash tools/profile.sh launches r3'
# Every pass first runs the same command WITHOUT ncu (it must exit 0), as /opt/skills/guides/B200_PROFILING.md asks.
# Output: gpurun_out/<tag>_*; summarise with `python tools/ncu_table.py gpurun_out/<tag>_..._raw.csv` and copy what
# is to be judged into profiles/.
#   launches   per-launch device time of two steady-state bench.py steps (cold cache, serialised: compare SHARES)
#   top        ncu --set full of k_linear_tc inside bench.py
#   kernel     ncu --set full of the launches of one kernel (regex) inside bench.py: profile.sh kernel <tag> <regex> [count]
#   linear_tc  ncu --set full of k_linear_tc / k_linear_tc_wgrad alone (tools/run_linear_tc_one.py, ..._wgrad_one.py)
#   fused      ncu --set full of the fused edge-conv kernels (tools/run_fused_kernels.py)
#   ops        ncu --set full of the a-row kernels at the odd widths (tools/run_ops_once.py)
#   rcd        counters of the classifier glue kernel (tools/ab_relu_concat_dropout.py)
set -x
what=${1:?what}; tag=${2:-prof}
mkdir -p gpurun_out
full() {  # full <name> <kernel regex> <count> <cmd...>
  local name=$1 rx=$2 n=$3; shift 3
  "$@" > gpurun_out/${tag}_${name}_plain.log 2>&1 || { tail -20 gpurun_out/${tag}_${name}_plain.log; exit 1; }
  timeout 900 ncu --set full --import-source on --clock-control none -k regex:"$rx" -c $n -o /tmp/${tag}_${name} --force-overwrite "$@" \
      > gpurun_out/${tag}_${name}_ncu.log 2>&1
  ncu -i /tmp/${tag}_${name}.ncu-rep --page raw --csv > gpurun_out/${tag}_ncu_full_${name}_raw.csv 2>/dev/null
  ls -la gpurun_out/${tag}_ncu_full_${name}_raw.csv
}
BENCH="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-ops --no-reference-cuda --no-other-configs"
case $what in
  launches)
    $BENCH > gpurun_out/${tag}_bench_for_ncu.json 2> gpurun_out/${tag}_bench_for_ncu.err || { tail -20 gpurun_out/${tag}_bench_for_ncu.err; exit 1; }
    timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 7000 --csv --log-file gpurun_out/${tag}_ncu_launches.csv $BENCH \
        > gpurun_out/${tag}_ncu_launches.log 2>&1
    wc -l gpurun_out/${tag}_ncu_launches.csv ;;
  top)       full top_kernel k_linear_tc 55 $BENCH ;;
  kernel)    full kernel_${3:?regex} "$3" ${4:-20} $BENCH ;;   # tools/profile.sh kernel <tag> <kernel regex> [launches]
  linear_tc) full linear_tc k_linear_tc 6 python tools/run_linear_tc_one.py
             full linear_tc_wgrad k_linear_tc_wgrad 2 python tools/run_linear_tc_wgrad_one.py ;;
  fused)     full fused "k_dense_mlp3|k_edge_combine|k_edge_xyz|k_pool_linear" 20 python tools/run_fused_kernels.py ;;
  ops)       full ops_odd_widths "k_flat|k_neighbor|k_seg|k_max|k_perm|k_row|k_segment" 60 python tools/run_ops_once.py \
                 --ops scatter,scatter_diff,gather,gather_diff,max_gather,max_scatter,permutate --channels 3,35,131 ;;
  rcd)       python tools/ab_relu_concat_dropout.py 515000 > gpurun_out/${tag}_rcd_ab.log 2>&1 || exit 1
             ncu --clock-control none -k regex:k_rcd_ --launch-count 4 --launch-skip 3 \
                 --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread \
                 --csv --log-file gpurun_out/${tag}_rcd_ncu.csv python tools/ab_relu_concat_dropout.py 515000 > gpurun_out/${tag}_rcd_ncu.log 2>&1 ;;
  *) echo "unknown pass $what"; exit 2 ;;
esac
true
