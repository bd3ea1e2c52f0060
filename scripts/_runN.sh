set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29812 bench.py --gpus 4 --steps 20 --warmup 5 --no-context > gpurun_out/r2t_bench_n4.json 2> gpurun_out/r2t_bench_n4.err
DDPM_EXCHANGE=two_stage timeout 300 $TR --master-port 29813 bench.py --gpus 4 --steps 20 --warmup 5 --no-context --no-parity-gate > gpurun_out/r2t_bench_n4_2stage.json 2> gpurun_out/r2t_bench_n4_2stage.err
timeout 600 $TR --master-port 29811 tests/multigpu_check.py > gpurun_out/r2t_mgc4.log 2>&1
timeout 120 $TR --master-port 29817 scripts/persist_timeline_ddp.py > gpurun_out/r2t_timeline_ddp4.txt 2>&1
