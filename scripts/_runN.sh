set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
DDPM_SMALL_EXCHANGE=one_shot timeout 300 $TR --master-port 29812 bench.py --gpus 8 --steps 20 --warmup 5 --no-context > gpurun_out/r2x_bench_n8_small1.json 2> gpurun_out/r2x_bench_n8_small1.err
timeout 300 $TR --master-port 29813 bench.py --gpus 8 --steps 20 --warmup 5 --no-context --no-parity-gate > gpurun_out/r2x_bench_n8_base.json 2> gpurun_out/r2x_bench_n8_base.err
DDPM_SMALL_EXCHANGE=one_shot timeout 300 $TR --master-port 29814 bench.py --gpus 8 --steps 200 --warmup 20 --no-context --no-parity-gate > gpurun_out/r2x_bench_n8_small1_200.json 2> gpurun_out/r2x_bench_n8_small1_200.err
