set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/r3a_pytest.log
timeout 300 python bench.py --steps 100 --warmup 20 --no-cpu-baseline --no-context > gpurun_out/r3a_bench.json 2> gpurun_out/r3a_bench.err
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-context --no-parity-gate > gpurun_out/r3a_bench20.json 2> gpurun_out/r3a_bench20.err
timeout 120 python scripts/persist_timeline.py 64 > gpurun_out/r3a_timeline.txt 2>&1
