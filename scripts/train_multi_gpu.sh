#!/bin/bash
# The reference's train_multi_gpu.sh (torch.distributed.launch --nproc_per_node=8) on the B200 engine.
# RANK/WORLD_SIZE/LOCAL_RANK/MASTER_* come from torchrun; nccl-mpich only reads them from the env.
NPROC="${NPROC:-8}"
python -m torch.distributed.run --nnodes=1 --nproc-per-node "${NPROC}" --master-addr 127.0.0.1 --master-port "${MASTER_PORT:-29500}" \
    "$(dirname "$0")/../mnist_pnetcdf_b200.py" --parallel --wireup_method nccl-mpich "$@"
