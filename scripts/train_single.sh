#!/bin/bash
# The reference's train_cpu.sh runs its single-process script; this is the single-GPU equivalent.
python "$(dirname "$0")/../mnist_pnetcdf_b200.py" "$@"
