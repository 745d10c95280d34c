# launch list of the bench command (one pass, time only); the same command must have exited 0 without ncu first
set -x
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/bench_for_ncu.json 2> gpurun_out/bench_for_ncu.err || { tail -20 gpurun_out/bench_for_ncu.err; exit 1; }
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 40000 --csv --log-file gpurun_out/launches_r1u.csv $CMD > gpurun_out/ncu_launches_r1u.log 2>&1
ls -la gpurun_out/launches_r1u.csv; tail -3 gpurun_out/ncu_launches_r1u.log
