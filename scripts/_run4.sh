set -x
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r2o_pytest.log
timeout 120 python scripts/persist_timeline.py 64 > gpurun_out/r2o_timeline.txt 2>&1
DDPM_LIB=/root/repo/pytorch_ddp_mnist_b200/lib_gt.so timeout 120 python scripts/persist_timeline_abs.py > gpurun_out/r2o_timeline_abs.txt 2>&1
timeout 300 python bench.py --steps 100 --warmup 20 --no-cpu-baseline > gpurun_out/r2o_bench.json 2> gpurun_out/r2o_bench.err
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-context --no-parity-gate > gpurun_out/r2o_bench20.json 2> gpurun_out/r2o_bench20.err
timeout 300 python bench.py --model json_convnet --steps 200 --warmup 10 --no-cpu-baseline > gpurun_out/r2o_convnet.json 2> gpurun_out/r2o_convnet.err
