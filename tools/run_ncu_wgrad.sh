set -x
CMD="python tools/dbg_wgrad.py"
$CMD || exit 1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_linear_tc_wgrad -c 2 -o /tmp/wg --force-overwrite $CMD > gpurun_out/ncu_wg.log 2>&1
ncu -i /tmp/wg.ncu-rep --page raw --csv > gpurun_out/ncu_wg_raw.csv 2>/dev/null
ncu -i /tmp/wg.ncu-rep --page source --csv > gpurun_out/ncu_wg_source.csv 2>/dev/null
ls -la gpurun_out/ncu_wg_raw.csv gpurun_out/ncu_wg_source.csv
