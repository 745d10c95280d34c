# final captures of the round with the current build: (1) plain bench, (2) ncu launch list of the same command,
# (3) ncu --set full of k_linear_tc and k_linear_tc_wgrad on the standalone drivers (small working sets: the capture
# inside the 24 GB bench took 19 GPU-minutes in r2g)
set -x
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/bench_for_ncu.json 2> gpurun_out/bench_for_ncu.err || { tail -20 gpurun_out/bench_for_ncu.err; exit 1; }
# -c 7000: the lazy-init forward (~800 launches) + five full steps (1055 launches each); the uncapped pass ran into its
# 600 s timeout after 9190 launches (11 GPU-minutes) — ncu costs ~65 ms per launch with a 24 GB context
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 7000 --csv --log-file gpurun_out/ncu_launches_r2k.csv $CMD > gpurun_out/ncu_launches_r2k.log 2>&1
wc -l gpurun_out/ncu_launches_r2k.csv
python tools/run_linear_tc_one.py || exit 1
timeout 300 ncu --set full --import-source on --clock-control none -k regex:k_linear_tc -c 6 -o /tmp/ltc --force-overwrite python tools/run_linear_tc_one.py > gpurun_out/ncu_ltc.log 2>&1
ncu -i /tmp/ltc.ncu-rep --page raw --csv > gpurun_out/ncu_ltc_r2k_raw.csv 2>/dev/null
python tools/dbg_wgrad.py || exit 1
timeout 300 ncu --set full --import-source on --clock-control none -k regex:k_linear_tc_wgrad -c 2 -o /tmp/wg --force-overwrite python tools/dbg_wgrad.py > gpurun_out/ncu_wg.log 2>&1
ncu -i /tmp/wg.ncu-rep --page raw --csv > gpurun_out/ncu_wg_r2k_raw.csv 2>/dev/null
ls -la gpurun_out/
true
