# ncu --set full of the dominant kernel inside the bench command: the 13 launches of one timed step
set -x
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/bench_for_ncu.json 2> gpurun_out/bench_for_ncu.err || { tail -20 gpurun_out/bench_for_ncu.err; exit 1; }
timeout 1200 ncu --set full --import-source on --clock-control none -k regex:k_dense_mlp3_bwd_mma --launch-skip 39 -c 13 -o /tmp/top --force-overwrite $CMD > gpurun_out/ncu_top_r1u.log 2>&1
ls -la /tmp/top.ncu-rep
ncu -i /tmp/top.ncu-rep --page raw --csv > gpurun_out/ncu_top_r1u_raw.csv 2>/dev/null
ncu -i /tmp/top.ncu-rep --page source --csv > gpurun_out/ncu_top_r1u_source.csv 2>/dev/null
ls -la gpurun_out/ncu_top_r1u_raw.csv gpurun_out/ncu_top_r1u_source.csv
[ $(stat -c %s /tmp/top.ncu-rep) -lt 30000000 ] && cp /tmp/top.ncu-rep gpurun_out/top_r1u.ncu-rep
true
