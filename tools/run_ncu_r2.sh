# round-1 final captures: (1) plain bench, (2) ncu launch list of the same command, (3) ncu --set full of the
# dominant kernel (k_linear_tc) inside the bench command
set -x
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/bench_for_ncu.json 2> gpurun_out/bench_for_ncu.err || { tail -20 gpurun_out/bench_for_ncu.err; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/ncu_launches_r2g.csv $CMD > gpurun_out/ncu_launches_r2g.log 2>&1
wc -l gpurun_out/ncu_launches_r2g.csv
timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_linear_tc -s 400 -c 55 -o /tmp/top --force-overwrite $CMD > gpurun_out/ncu_top_r2g.log 2>&1
ls -la /tmp/top.ncu-rep
ncu -i /tmp/top.ncu-rep --page raw --csv > gpurun_out/ncu_top_r2g_raw.csv 2>/dev/null
ls -la gpurun_out/ncu_top_r2g_raw.csv
true
