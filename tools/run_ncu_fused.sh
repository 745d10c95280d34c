set -x
CMD="python tools/run_fused_kernels.py"
$CMD > gpurun_out/fused_plain.log 2>&1 || { tail -20 gpurun_out/fused_plain.log; exit 1; }
cat gpurun_out/fused_plain.log
timeout 800 ncu --set full --import-source on --clock-control none -k regex:"k_dense_mlp3|k_edge_combine|k_edge_xyz|k_pool_linear" -c 20 -o /tmp/fused --force-overwrite $CMD > gpurun_out/ncu_fused.log 2>&1
ls -la /tmp/fused.ncu-rep
ncu -i /tmp/fused.ncu-rep --page raw --csv > gpurun_out/ncu_fused_raw.csv 2>/dev/null
ncu -i /tmp/fused.ncu-rep --page details --csv > gpurun_out/ncu_fused_details.csv 2>/dev/null
ls -la gpurun_out/ncu_fused_raw.csv gpurun_out/ncu_fused_details.csv
[ $(stat -c %s /tmp/fused.ncu-rep) -lt 40000000 ] && cp /tmp/fused.ncu-rep gpurun_out/fused_r1o.ncu-rep
