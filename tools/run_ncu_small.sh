set -x
CMD="python bench_ops.py --channels 3,32,35 --iters 1 --no-ref --no-cpu --only neighbor_scatter,sum_feat_gather,max_feat_gather,neighbor_gather"
$CMD > gpurun_out/ops_small_plain.log 2>&1 || { tail -20 gpurun_out/ops_small_plain.log; exit 1; }
tail -30 gpurun_out/ops_small_plain.log
timeout 800 ncu --set full --clock-control none -k regex:"k_flat_rows|k_segment_sum|k_max_gather|k_neighbor_gather|k_neighbor_scatter" -c 80 -o /tmp/small_c --force-overwrite $CMD > gpurun_out/ncu_small.log 2>&1
ls -la /tmp/small_c.ncu-rep
ncu -i /tmp/small_c.ncu-rep --page raw --csv > gpurun_out/ncu_small_raw.csv 2>/dev/null
ls -la gpurun_out/ncu_small_raw.csv
python bench_ops.py --no-cpu --out gpurun_out/ops_r1k.json > gpurun_out/ops_r1k.log 2>&1; tail -5 gpurun_out/ops_r1k.log
