set -x
CMD="python tools/run_linear_tc_one.py"
$CMD || exit 1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_linear_tc -c 6 -o /tmp/ltc --force-overwrite $CMD > gpurun_out/ncu_ltc.log 2>&1
ncu -i /tmp/ltc.ncu-rep --page raw --csv > gpurun_out/ncu_ltc_raw.csv 2>/dev/null
ncu -i /tmp/ltc.ncu-rep --page source --csv > gpurun_out/ncu_ltc_source.csv 2>/dev/null
ls -la gpurun_out/ncu_ltc_raw.csv gpurun_out/ncu_ltc_source.csv
