"""Process-group wire-up: the reference's `class distributed` (mp:51-272) re-done as a table.

The reference discovers MASTER_ADDR / MASTER_PORT / WORLD_SIZE / RANK from whichever launcher is
in use and then calls `dist.init_process_group(backend, init_method='env://')`.  Same method names
and the same precedence here; the two latent bugs of the reference are not reproduced
(`os.environ("PMIX_SERVER_URI2")` called instead of indexed, mp:134; `ntasks_per_node * num_nodes`
multiplying two strings, mp:118).  `mpich` maps to backend "mpi", which needs a torch built with
MPI (this image has none): it raises with that explanation instead of failing inside torch.
"""
from __future__ import annotations

import os
from typing import Optional

import torch
import torch.distributed as dist

METHODS = ("nccl-slurm", "nccl-openmpi", "nccl-mpich", "gloo", "mpich")
_BACKEND = {"nccl-slurm": "nccl", "nccl-openmpi": "nccl", "nccl-mpich": "nccl", "gloo": "gloo", "mpich": "mpi"}


def _mpi_world():
    try:
        from mpi4py import MPI
        if MPI.Is_initialized():
            return MPI.COMM_WORLD.Get_size(), MPI.COMM_WORLD.Get_rank()
    except Exception:
        pass
    return None


def _first(*names: str) -> Optional[str]:
    for n in names:
        if n in os.environ:
            return os.environ[n]
    return None


def _pmix_host() -> Optional[str]:
    uri = os.environ.get("PMIX_SERVER_URI2")
    if not uri or "//" not in uri:
        return None
    return uri.split("//")[1].split(":")[0]


def discover_env(method: str) -> None:
    """Fill MASTER_ADDR/MASTER_PORT/WORLD_SIZE/RANK in os.environ the way the reference's branch
    for `method` does; explicit values already in the environment always win."""
    if method not in METHODS:
        raise NotImplementedError(method)
    env = os.environ
    mpi = _mpi_world()
    if method == "nccl-slurm":                                                 # mp:84-129
        if "MASTER_ADDR" not in env:
            addr = _first("SLURM_LAUNCH_NODE_IPADDR")
            if addr is None:
                raise Exception("Error: nccl-slurm - SLURM_LAUNCH_NODE_IPADDR is not set")
            env["MASTER_ADDR"] = addr
        env.setdefault("MASTER_PORT", _first("SLURM_SRUN_COMM_PORT") or "29500")
        if "WORLD_SIZE" not in env:
            ws = _first("SLURM_NTASKS")
            if ws is None:
                nodes = _first("SLURM_JOB_NUM_NODES")
                if nodes is None:
                    raise Exception("Error: nccl-slurm - SLURM_JOB_NUM_NODES is not set")
                per = _first("SLURM_NTASKS_PER_NODE", "SLURM_TASKS_PER_NODE")
                if per is None:
                    raise Exception("Error: nccl-slurm - SLURM_(N)TASKS_PER_NODE is not set")
                ws = str(int(per.split("(")[0]) * int(nodes))
            env["WORLD_SIZE"] = str(ws)
        if "RANK" not in env:
            r = _first("SLURM_PROCID")
            if r is None:
                raise Exception("Error: nccl-slurm - SLURM_PROCID is not set")
            env["RANK"] = r
    elif method == "nccl-openmpi":                                             # mp:131-153
        if "MASTER_ADDR" not in env:
            host = _pmix_host()
            if host is None:
                raise Exception("Error: nccl-openmpi - PMIX_SERVER_URI2 is not set")
            env["MASTER_ADDR"] = host
        env.setdefault("MASTER_PORT", "29500")
        for key, src in (("WORLD_SIZE", "OMPI_COMM_WORLD_SIZE"), ("RANK", "OMPI_COMM_WORLD_RANK")):
            if key not in env:
                if src not in env:
                    raise Exception(f"Error: nccl-openmpi - {src} is not set")
                env[key] = env[src]
    else:                                                                      # nccl-mpich / mpich / gloo
        if "MASTER_ADDR" not in env:
            env["MASTER_ADDR"] = (_pmix_host() if method == "gloo" else None) or "localhost"   # mp:214-221
        env.setdefault("MASTER_PORT", "29500")
        if "WORLD_SIZE" not in env:
            names = ("OMPI_COMM_WORLD_SIZE", "PMI_SIZE") if method == "gloo" else ("PMI_SIZE",)
            env["WORLD_SIZE"] = str(_first(*names) or (mpi[0] if mpi else 1))
        if "RANK" not in env:
            names = ("OMPI_COMM_WORLD_RANK", "PMI_RANK") if method == "gloo" else ("PMI_RANK",)
            env["RANK"] = str(_first(*names) or (mpi[1] if mpi else 0))


class distributed:
    """Same surface as the reference class: `distributed(method)`, `get_size()`, `get_rank()`,
    `get_local_rank()`, `barrier()`, `reduceMAX()`, `finalize()`."""

    def __init__(self, method: str, backend: str = None):
        """`backend` overrides the transport the reference ties to `method` (train.py runs the `gloo` wire-up -
        its OMPI_COMM_WORLD_* / PMIx / PMI discovery, mp:213-254 - over NCCL, because the data lives on GPUs)."""
        discover_env(method)
        backend = backend or _BACKEND[method]
        if backend == "mpi" and not dist.is_mpi_available():
            raise NotImplementedError("wireup_method 'mpich' needs a torch build with the MPI backend; "
                                      "use 'gloo' or one of the nccl-* methods")
        kw = {}
        if backend == "nccl":
            local = int(os.environ.get("LOCAL_RANK", int(os.environ["RANK"]) % max(torch.cuda.device_count(), 1)))
            torch.cuda.set_device(local)
            kw["device_id"] = torch.device("cuda", local)
        dist.init_process_group(backend=backend, init_method="env://", **kw)

    def get_size(self) -> int:
        return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1

    def get_rank(self) -> int:
        return dist.get_rank() if dist.is_available() and dist.is_initialized() else 0

    def get_local_rank(self) -> int:
        if not (dist.is_available() and dist.is_initialized()):
            return 0
        if torch.cuda.is_available():
            return int(os.environ.get("LOCAL_RANK", dist.get_rank() % torch.cuda.device_count()))
        return -1

    def reduceMAX(self, src, root):
        """The reference does this over MPI (mp:259-265, never called); here over the process group."""
        t = torch.as_tensor(src, dtype=torch.float64).clone()
        if dist.is_initialized():
            if dist.get_backend() == "nccl":
                t = t.cuda()
            dist.reduce(t, root, op=dist.ReduceOp.MAX)
        return t.cpu().numpy()

    def barrier(self) -> None:
        if dist.is_initialized():
            dist.barrier()

    def finalize(self) -> None:
        if dist.is_initialized():
            dist.destroy_process_group()
