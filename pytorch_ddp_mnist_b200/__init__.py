"""B200-native data-parallel MNIST training step (drop-in for the hot path of
Jonathanlyj/pytorch_ddp_mnist's `mnist_pnetcdf_cpu_mp.py`).

Host-only modules (`cdf5`, `synth`, `compat`) import without the CUDA library; everything else
binds `libddpmnist_b200.so` on first use and raises if it is missing (no CPU fallback).
"""
import importlib

__version__ = "0.1.0"
_LAZY = {"ops", "engine", "dataset", "model", "train", "wireup", "_lib"}


def __getattr__(name):
    if name in _LAZY:
        return importlib.import_module(f"{__name__}.{name}")
    raise AttributeError(name)
