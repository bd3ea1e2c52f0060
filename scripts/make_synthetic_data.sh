#!/bin/bash
# Writes mnist_train_images.nc / mnist_test_images.nc (seeded synthetic, CDF-5) into the cwd.
python -m pytorch_ddp_mnist_b200.synth --root "${1:-.}"
