set -x
python scripts/run_large_once.py 4096 > gpurun_out/r2y_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r2y_large_launches.csv python scripts/run_large_once.py 4096 > gpurun_out/r2y_ncu1.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"fc1_fwd_tc|dw1_tc|rows_gemm|grads_kernel" -s 4 -c 4 -o gpurun_out/r2y_large_prof -f python scripts/run_large_once.py 4096 > gpurun_out/r2y_ncu2.log 2>&1
