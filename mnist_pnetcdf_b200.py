#!/usr/bin/env python
"""Drop-in for `python mnist_pnetcdf_cpu_mp.py [flags]` on B200 GPUs: same flags, same printed
lines, same `model.pt` (see pytorch_ddp_mnist_b200/train.py).

    python mnist_pnetcdf_b200.py --batch_size 64 --n_epochs 1
    torchrun --standalone --nproc-per-node 8 mnist_pnetcdf_b200.py --parallel --wireup_method nccl-mpich --batch_size 64
"""
from pytorch_ddp_mnist_b200.train import run

if __name__ == '__main__':
    run()
