N=$1
set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29811 tests/multigpu_check.py > gpurun_out/r2j_mgc$N.log 2>&1
timeout 300 $TR --master-port 29812 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2j_bench_n$N.json 2> gpurun_out/r2j_bench_n$N.err
timeout 300 $TR --master-port 29813 bench.py --gpus $N --steps 200 --warmup 20 --no-context --no-parity-gate > gpurun_out/r2j_bench_n${N}_200.json 2> gpurun_out/r2j_bench_n${N}_200.err
DDPM_ACT_EXCHANGE=1 timeout 300 $TR --master-port 29815 bench.py --gpus $N --steps 200 --warmup 20 --no-context --no-parity-gate > gpurun_out/r2j_bench_n${N}_act.json 2> gpurun_out/r2j_bench_n${N}_act.err
DDPM_NO_ACT_EXCHANGE=1 timeout 120 $TR --master-port 29816 scripts/persist_timeline_ddp.py > gpurun_out/r2j_timeline_ddp${N}_bucket.txt 2>&1
timeout 300 $TR --master-port 29818 bench.py --impl reference --gpus $N --steps 20 --warmup 5 > gpurun_out/r2j_ref_n$N.json 2> gpurun_out/r2j_ref_n$N.err
