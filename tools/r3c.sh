mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r3c_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r3c_pytest.log; tail -4 gpurun_out/r3c_pytest.log
python bench_ops.py --dtypes f32,bf16 --out gpurun_out/r3c_ops.json > gpurun_out/r3c_ops.log 2>&1; echo "ops rc=$?"; tail -2 gpurun_out/r3c_ops.log
python tools/run_ops_once.py --ops scatter,scatter_diff,gather,gather_diff,max_gather,permutate --channels 3,35,131 > gpurun_out/r3c_once.log 2>&1 && \
ncu --set full --clock-control none -k regex:'k_flat_rows|k_neighbor_scatter|k_neighbor_gather|k_seg_stream' -c 40 -o /tmp/r3c_prof python tools/run_ops_once.py --ops scatter,scatter_diff,gather,gather_diff,max_gather,permutate --channels 3,35,131 > gpurun_out/r3c_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/r3c_ncu.log
ncu -i /tmp/r3c_prof.ncu-rep --page raw --csv > gpurun_out/r3c_ncu_raw.csv 2>/dev/null; ls -la gpurun_out
