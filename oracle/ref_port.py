"""CPU port of the reference's training path, used ONLY as the timed CPU baseline
(`bench.py` cpu_baseline leg and `bench.py --impl reference`) and as a checker in tests.

TEST/BENCH INFRASTRUCTURE — not a product path; nothing under pytorch_ddp_mnist_b200/ imports it.

Why a port: `/root/reference` is Python and cannot travel to the GPU box.  This file restates
`mnist_pnetcdf_cpu_mp.py` with the SAME libraries and the same cost structure — per-sample
positioned reads through the `pncpy` API (mp:39-49), torchvision `ToTensor`+`Normalize`
(mp:374-377), `DataLoader(num_workers=0)` + `DistributedSampler(seed=42)` (mp:390-407),
`nn.Sequential` MLP (mp:412-422), `DistributedDataParallel` over gloo (mp:439), `optim.SGD(lr=0.01)`
and `nn.CrossEntropyLoss` (mp:443-444), and the loop body with its per-step `.item()` (mp:453-466).
`tests/test_ref_port.py` checks that it reproduces the losses the real reference produced
(tests/golden/traj_ws1.npz) so the baseline being timed is the reference's arithmetic.
"""
from __future__ import annotations

import os
import sys
import time
from typing import Dict, List, Optional

import numpy as np

_REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (os.path.join(_REPO, "pytorch_ddp_mnist_b200", "compat"), _REPO):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from torch import nn, optim  # noqa: E402
from torch.nn.parallel.distributed import DistributedDataParallel  # noqa: E402
from torch.utils.data import DataLoader, Dataset, DistributedSampler  # noqa: E402


class PortMNISTNetCDF(Dataset):
    """mp:18-49 — one image read + one label read + transform per item."""

    def __init__(self, root_dir, is_train=True, transforms=None, comm=None):
        import pncpy
        self.transforms = transforms
        nc_path = os.path.join(root_dir, 'mnist_{}_images.nc'.format('train' if is_train else 'test'))
        self.nc = pncpy.File(nc_path, 'r', comm=comm)
        self.nc.begin_indep()

    def __len__(self):
        return self.nc.variables['images'].shape[0]

    def __getitem__(self, index):
        image = np.array(self.nc.variables['images'][index])
        buff = np.empty((), np.uint8)
        self.nc.variables['labels'].get_var(buff, index=(index,))
        if self.transforms:
            image = self.transforms(image)
        return image, buff


def port_create_model():
    """mp:412-422."""
    return nn.Sequential(nn.Linear(28 * 28, 128), nn.ReLU(), nn.Dropout(0.2), nn.Linear(128, 128), nn.ReLU(),
                         nn.Linear(128, 10, bias=False))


def port_train_loader(root_dir: str, batch_size: int, rank: int, world_size: int):
    """mp:370-401."""
    from torchvision import transforms
    transform = transforms.Compose([transforms.ToTensor(), transforms.Normalize((0.1307,), (0.3081,))])
    ds = PortMNISTNetCDF(root_dir, is_train=True, transforms=transform)
    sampler = DistributedSampler(ds, num_replicas=world_size, rank=rank, shuffle=True, seed=42)
    return DataLoader(ds, batch_size=batch_size, shuffle=False, num_workers=0, sampler=sampler, pin_memory=False)


def port_train(root_dir: str, batch_size: int, steps: int, warmup: int, rank: int = 0, world_size: int = 1,
               model_seed: Optional[int] = 0, mask_feed=None, record_losses: bool = False) -> Dict:
    """The loop body of mp:447-466 for `warmup + steps` batches of epoch 0; times the last `steps`.
    Needs an initialised process group when world_size > 1 (gloo)."""
    if model_seed is not None:
        torch.manual_seed(model_seed + rank)
    loader = port_train_loader(root_dir, batch_size, rank, world_size)
    device = torch.device("cpu")
    model = port_create_model().to(device)
    if dist.is_initialized():
        model = DistributedDataParallel(model)
    optimizer = optim.SGD(model.parameters(), lr=0.01)
    loss = nn.CrossEntropyLoss()
    model.train()
    loader.sampler.set_epoch(0)
    losses: List[float] = []
    epoch_loss, n_img, t0, t_load = 0.0, 0, None, 0.0
    it = iter(loader)
    for t in range(warmup + steps):
        if t == warmup:
            if dist.is_initialized():
                dist.barrier()
            t0 = time.perf_counter()
        tl = time.perf_counter()
        try:
            x, y = next(it)
        except StopIteration:
            it = iter(loader)
            x, y = next(it)
        if t >= warmup:
            t_load += time.perf_counter() - tl
        x = x.to(device, non_blocking=True)
        y = y.to(device, non_blocking=True)
        x = x.view(x.shape[0], -1)
        optimizer.zero_grad()
        y_hat = model(x)
        batch_loss = loss(y_hat, y)
        batch_loss.backward()
        optimizer.step()
        batch_loss_scalar = batch_loss.item()
        epoch_loss += batch_loss_scalar / x.shape[0]
        if record_losses:
            losses.append(batch_loss_scalar)
        if t >= warmup:
            n_img += x.shape[0]
    if dist.is_initialized():
        dist.barrier()
    dt = time.perf_counter() - t0
    m = model.module if isinstance(model, DistributedDataParallel) else model
    return {"seconds": dt, "images": n_img, "loader_seconds": t_load, "losses": losses,
            "state": {k: v.detach().clone() for k, v in m.state_dict().items()}}


# Variables a launcher (torchrun / torch.distributed.launch) exports into its workers.  A process spawned from such a
# worker inherits them; with TORCHELASTIC_USE_AGENT_STORE set, torch's rendezvous makes even rank 0 a *client* of the
# launcher's store, so a private group built by env:// or tcp:// rendezvous waits for a server nobody starts.
_LAUNCHER_ENV = ("TORCHELASTIC_", "TORCH_NCCL_", "NCCL_ASYNC", "ROLE_", "GROUP_", "LOCAL_RANK", "LOCAL_WORLD_SIZE",
                 "RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT", "OMP_NUM_THREADS")


def scrub_launcher_env(env=None):
    """Remove every launcher-exported variable from `env` (default os.environ); returns the removed names."""
    env = os.environ if env is None else env
    gone = [k for k in list(env) if k.startswith(_LAUNCHER_ENV)]
    for k in gone:
        del env[k]
    return gone


def _worker(rank, world_size, store_path, root_dir, batch_size, steps, warmup, threads, q):
    try:
        scrub_launcher_env()
        torch.set_num_threads(max(1, threads))
        # private rendezvous through a file: no TCP port to collide with, nothing read from the environment
        store = dist.FileStore(store_path, world_size)
        import datetime
        dist.init_process_group("gloo", store=store, rank=rank, world_size=world_size,
                                timeout=datetime.timedelta(seconds=120))
        r = port_train(root_dir, batch_size, steps, warmup, rank, world_size)
        dist.destroy_process_group()
        q.put((rank, r["seconds"], r["images"], r["loader_seconds"], None))
    except BaseException as e:                        # the parent must hear about it instead of waiting
        q.put((rank, 0.0, 0, 0.0, repr(e)))
        raise


def time_port_gloo(root_dir: str, batch_size: int, steps: int, warmup: int, world_size: int = 2,
                   port: int = 0, timeout_s: float = 600.0) -> Dict:
    """BASELINE.json configs[0]: `--parallel --wireup_method gloo --batch_size 64`, world_size
    processes x (cores // world_size) threads; returns global images/s over the timed steps
    (max over ranks of the wall time, like the GPU arm).  Safe to call from a torchrun worker: the
    children drop the launcher's environment and rendezvous through a private FileStore (`port` is
    unused, kept for callers of the old signature).  Raises RuntimeError (never hangs) if a worker
    dies or the whole run exceeds `timeout_s`."""
    import queue as _queue
    import tempfile
    import torch.multiprocessing as mp
    cores = os.cpu_count() or 1
    threads = max(1, cores // world_size)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    store_dir = tempfile.mkdtemp(prefix="refport_store_")
    store_path = os.path.join(store_dir, "store")
    ps = [ctx.Process(target=_worker, args=(r, world_size, store_path, root_dir, batch_size, steps, warmup, threads, q),
                      daemon=True) for r in range(world_size)]
    saved = dict(os.environ)                          # children are spawned with a launcher-free environment
    scrub_launcher_env()
    try:
        for p in ps:
            p.start()
    finally:
        os.environ.update({k: v for k, v in saved.items() if k not in os.environ})
    res, deadline, err = [], time.monotonic() + timeout_s, None
    try:
        while len(res) < world_size and err is None:
            try:
                item = q.get(timeout=1.0)
                if item[4] is not None:
                    err = f"rank {item[0]} failed: {item[4]}"
                res.append(item)
            except _queue.Empty:
                dead = [p for p in ps if p.exitcode not in (None, 0)]
                if dead:
                    err = f"worker exited with code {dead[0].exitcode}"
                elif time.monotonic() > deadline:
                    err = f"timed out after {timeout_s:.0f} s"
    finally:
        for p in ps:
            p.join(timeout=5.0 if err is None else 0.5)
            if p.is_alive():
                p.terminate()
                p.join(timeout=5.0)
        import shutil
        shutil.rmtree(store_dir, ignore_errors=True)
    if err is not None:
        raise RuntimeError(f"time_port_gloo(world_size={world_size}): {err}")
    secs = max(r[1] for r in res)
    imgs = sum(r[2] for r in res)
    return {"images_per_s": imgs / secs, "seconds": secs, "images": imgs, "world_size": world_size,
            "threads_per_rank": threads, "cores": min(cores, threads * world_size),
            "loader_fraction": max(r[3] for r in res) / secs}
