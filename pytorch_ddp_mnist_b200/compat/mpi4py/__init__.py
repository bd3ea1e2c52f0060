"""`mpi4py` stand-in (no MPI in this image).

The reference only needs `from mpi4py import MPI` to hand `MPI.COMM_WORLD` to PnetCDF
(`mnist_pnetcdf_cpu_mp.py:11,493,504`) and, in the env-discovery fallbacks, `MPI.Is_initialized()`
(`mp:165,194,234,247`).  Under a non-MPI launcher every process is a singleton MPI world, which is
what this models: rank 0 of 1.
"""
from . import MPI  # noqa: F401

__version__ = "0.0-b200-compat"
