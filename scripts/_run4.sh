set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r2u_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2u_smoke.log 2>&1
timeout 600 python bench.py > gpurun_out/r2u_bench_n1_default.json 2> gpurun_out/r2u_bench_n1_default.err
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2u_bench_n1_20.json 2> gpurun_out/r2u_bench_n1_20.err
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2u_ref_n1.json 2> gpurun_out/r2u_ref_n1.err
timeout 120 python scripts/persist_timeline.py 64 > gpurun_out/r2u_timeline.txt 2>&1
timeout 200 python scripts/time_host_fed.py 64 4000 > gpurun_out/r2u_hostfed.txt 2>&1
python scripts/ncu_persist.py > gpurun_out/r2u_plain.log 2>&1 && {
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/r2u_launches.csv python scripts/ncu_persist.py > gpurun_out/r2u_ncu1.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:train_persist -s 2 -c 3 -o gpurun_out/r2u_prof -f python scripts/ncu_persist.py > gpurun_out/r2u_ncu2.log 2>&1
}
