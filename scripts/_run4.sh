set -x
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/r2s_pytest.log
timeout 120 python scripts/persist_timeline.py 64 > gpurun_out/r2s_timeline.txt 2>&1
