#!/bin/bash
# The reference's train_cpu_mp.csh is `mpiexec -n 4 python3 mnist_pnetcdf_cpu_mp.py --parallel --wireup_method mpich`.
# There is no MPI on the B200 box: same 4 ranks through torchrun, one GPU each.
NPROC="${NPROC:-4}" exec "$(dirname "$0")/train_multi_gpu.sh" "$@"
