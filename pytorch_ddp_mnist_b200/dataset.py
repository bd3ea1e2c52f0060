"""Dataset side of the path: the netCDF file goes to HBM once, batches are formed on the device.

Reference behaviour being replaced (`mnist_pnetcdf_cpu_mp.py`, "mp:"):
  * `MNISTNetCDF` (mp:18-49) opens `root_dir/mnist_{train,test}_images.nc`, and every
    `__getitem__` does one 784-byte read, one 1-byte read and three tiny tensor ops;
  * `DataLoader(num_workers=0)` (mp:396-407) runs that B times per batch on the training thread.
Here the file is read ONCE with large positioned reads (the C reader, `ddpm_cdf5_read_rows`) into
pinned staging buffers that alternate with `cudaMemcpyAsync`, so file reads overlap the copies; with
several ranks each rank reads only its contiguous 1/ws slab and an NCCL all-gather replicates the
uint8 dataset (47 MB, or 4.7 GB for the 6M-sample file — small against 180 GB of HBM).  A rank's
shard is `perm[rank::ws]`, scattered over the whole file, which is why every GPU keeps all of it.
Per step the only loader work left is the gather+decode kernel (or nothing, when the decode is
fused into fc1's operand load).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Iterator, Optional, Tuple

import numpy as np
import torch
from torch.utils.data import DataLoader, Dataset, DistributedSampler

from . import ops
from ._lib import DDPMError, check, lib

IMG_BYTES = 784


class CDF5File:
    """Thin RAII wrapper over the C reader."""

    def __init__(self, path: str):
        self.path = path
        self._h = C.c_void_p()
        check(lib.ddpm_cdf5_open(os.fsencode(path), C.byref(self._h)))

    def inq(self, name: str) -> Tuple[Tuple[int, ...], int, int, int]:
        nd, dims, t, es, beg = C.c_int(), (C.c_int64 * 8)(), C.c_int(), C.c_int(), C.c_int64()
        check(lib.ddpm_cdf5_inq_var(self._h, name.encode(), C.byref(nd), dims, C.byref(t), C.byref(es),
                                    C.byref(beg)))
        return tuple(dims[:nd.value]), t.value, es.value, beg.value

    def read_rows(self, name: str, row0: int, nrows: int, dst: torch.Tensor) -> None:
        check(lib.ddpm_cdf5_read_rows(self._h, name.encode(), row0, nrows, dst.data_ptr()))

    def close(self) -> None:
        if self._h:
            lib.ddpm_cdf5_close(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_PINNED = {}


def _pinned_pair(slot: int, rows: int):
    """Two page-locked staging buffers per reader thread, kept for the life of the process (cudaHostAlloc is slow)."""
    key = (slot, rows)
    if key not in _PINNED:
        _PINNED[key] = [torch.empty((rows, IMG_BYTES), dtype=torch.uint8).pin_memory() for _ in range(2)]
    return _PINNED[key]


class ResidentDataset:
    """u8 images [N,784] and labels [N] resident on one GPU."""

    def __init__(self, images: torch.Tensor, labels: torch.Tensor):
        if images.dtype != torch.uint8 or labels.dtype != torch.uint8:
            raise TypeError("images and labels must be uint8")
        self.images = images.reshape(images.shape[0], IMG_BYTES).contiguous()
        self.labels = labels.contiguous()
        if self.images.shape[0] != self.labels.shape[0]:
            raise ValueError("images/labels length mismatch")

    def __len__(self) -> int:
        return self.images.shape[0]

    @property
    def device(self) -> torch.device:
        return self.images.device

    @classmethod
    def from_arrays(cls, images: np.ndarray, labels: np.ndarray, device) -> "ResidentDataset":
        return cls(torch.from_numpy(np.ascontiguousarray(images).reshape(len(labels), IMG_BYTES)).to(device),
                   torch.from_numpy(np.ascontiguousarray(labels)).to(device))

    @classmethod
    def from_file(cls, path: str, device, rank: int = 0, world_size: int = 1, group=None,
                  chunk_rows: int = 1 << 16, stats: Optional[dict] = None) -> "ResidentDataset":
        """Read a CDF-5 MNIST file into HBM.  With world_size > 1 every rank reads rows
        [rank*per, (rank+1)*per) and the slabs are all-gathered on `group` (NCCL).
        `stats` (optional dict) receives the host-clock seconds of the two phases: `read_copy_s` (pread into the
        pinned staging buffers overlapped with the H2D copies, until the last copy has landed) and `allgather_s`."""
        import time as _time
        _t0 = _time.perf_counter()
        f = CDF5File(path)
        try:
            shape, nc_type, es, _ = f.inq("images")
            lshape, ltype, les, _ = f.inq("labels")
            if len(shape) != 3 or shape[1] * shape[2] != IMG_BYTES or es != 1 or les != 1 or lshape != (shape[0],):
                raise DDPMError(f"{path}: expected images u8[N,28,28] and labels u8[N], got {shape} / {lshape}")
            n = shape[0]
            per = -(-n // world_size)
            r0, r1 = min(rank * per, n), min((rank + 1) * per, n)
            dev = torch.device(device)
            gather = world_size > 1
            images = torch.empty((per * world_size if gather else n, IMG_BYTES), dtype=torch.uint8, device=dev)
            labels = torch.empty((per * world_size if gather else n,), dtype=torch.uint8, device=dev)
            base = rank * per if gather else 0
            # `nthreads` reader threads (DDPM_LOADER_THREADS; default 8, fewer on small hosts), each with two pinned staging buffers and its
            # own copy stream: pread (positioned, thread-safe; the C call drops the GIL) into one buffer while the other
            # is in flight to the GPU.  Measured on the 6 M-sample file (4.7 GB, page cache warm): 1 thread 5.9 GB/s,
            # 4 threads 22.9 GB/s, 8 threads 34 GB/s into HBM (profiles/r02_scaled6m_*.json).
            import threading
            nthreads = max(1, int(os.environ.get("DDPM_LOADER_THREADS",
                                                 str(min(8, max(2, (os.cpu_count() or 8) // max(world_size, 1)))))))
            chunk_rows = min(chunk_rows, 1 << 14)
            starts = list(range(r0, r1, chunk_rows))
            errors = []

            def reader(tix: int) -> None:
                try:
                    torch.cuda.set_device(dev)
                    st = torch.cuda.Stream(device=dev)
                    bufs = _pinned_pair(tix, chunk_rows)
                    evs = [torch.cuda.Event(), torch.cuda.Event()]
                    for i, s0 in enumerate(starts[tix::nthreads]):
                        m = min(chunk_rows, r1 - s0)
                        b = i & 1
                        if i >= 2:
                            evs[b].synchronize()
                        f.read_rows("images", s0, m, bufs[b])
                        with torch.cuda.stream(st):
                            images[base + s0 - r0: base + s0 - r0 + m].copy_(bufs[b][:m], non_blocking=True)
                            evs[b].record(st)
                    st.synchronize()
                except BaseException as e:               # surfaced by the caller
                    errors.append(e)

            threads = [threading.Thread(target=reader, args=(t,)) for t in range(min(nthreads, max(len(starts), 1)))]
            for t in threads:
                t.start()
            for t in threads:
                t.join()
            if errors:
                raise errors[0]
            copy_stream = torch.cuda.Stream(device=dev)
            lab = torch.empty(max(r1 - r0, 1), dtype=torch.uint8).pin_memory()
            f.read_rows("labels", r0, r1 - r0, lab)
            with torch.cuda.stream(copy_stream):
                labels[base: base + r1 - r0].copy_(lab[:r1 - r0], non_blocking=True)
            copy_stream.synchronize()
        finally:
            f.close()
        _t1 = _time.perf_counter()
        if gather:
            import torch.distributed as dist
            dist.all_gather_into_tensor(images.view(-1), images[base: base + per].reshape(-1).clone(), group=group)
            dist.all_gather_into_tensor(labels, labels[base: base + per].clone(), group=group)
            images, labels = images[:n], labels[:n]
            torch.cuda.synchronize(dev)
        if stats is not None:
            stats["read_copy_s"] = _t1 - _t0
            stats["allgather_s"] = (_time.perf_counter() - _t1) if gather else 0.0
            stats["rows_read"] = r1 - r0
        return cls(images, labels)


class MNISTNetCDF(Dataset):
    """Drop-in for the reference class of the same name (mp:18-49): same constructor, `len(ds)`,
    and `ds[i] -> (image, label)`.  `transforms` is accepted for signature compatibility; the
    decode the reference composes there (ToTensor + Normalize, mp:374-377) is what the gather
    kernel computes, so it is applied iff `transforms` is not None.  Items are produced on the GPU
    (image f32[1,28,28] CUDA tensor, label 0-d np.uint8 like the reference's `buff`)."""

    def __init__(self, root_dir, is_train=True, transforms=None, comm=None, device=None,
                 rank: int = 0, world_size: int = 1, group=None):
        self.transforms = transforms
        print('=> Reading NetCDF File...')
        nc_path = os.path.join(root_dir, 'mnist_{}_images.nc'.format('train' if is_train else 'test'))
        dev = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.resident = ResidentDataset.from_file(nc_path, dev, rank, world_size, group)
        self.comm = comm
        print('=> Dataset created, image nc file is : {}'.format(nc_path))

    def __len__(self):
        return len(self.resident)

    def __getitem__(self, index):
        r = self.resident
        n = len(r)
        if not -n <= index < n:
            raise IndexError(index)
        idx = torch.tensor([index % n], dtype=torch.int64, device=r.device)
        if self.transforms:
            x, y = ops.gather_decode(r.images, r.labels, idx)
            image = x.view(1, 28, 28)
        else:
            image = r.images[index % n].view(28, 28)
            y = r.labels[idx]
        return image, np.array(y.cpu().numpy()[0], dtype=np.uint8)


class _IndexOnly(Dataset):
    """Stands in for the dataset inside torch's DataLoader so that samplers, batching, epoch
    reseeding and global-RNG consumption are exactly the reference's; items are just indices."""

    def __init__(self, n: int):
        self.n = n

    def __len__(self):
        return self.n

    def __getitem__(self, i):
        return i


class DeviceBatchLoader:
    """Iterates like the reference's DataLoader (mp:396-407) — same sampler objects, same batch
    boundaries, same short last batch — but each batch is formed by the gather+decode kernel:
    yields (x f32[B,1,28,28], y u8[B]) CUDA tensors.  `.sampler` is exposed for `set_epoch`."""

    def __init__(self, dataset, batch_size: int, shuffle: bool = False, sampler=None,
                 num_workers: int = 0, pin_memory: bool = True, decode: bool = True):
        self.resident = dataset.resident if isinstance(dataset, MNISTNetCDF) else dataset
        self.batch_size = batch_size
        self._index_loader = DataLoader(_IndexOnly(len(self.resident)), batch_size=batch_size,
                                        shuffle=shuffle, sampler=sampler, num_workers=0)
        self.decode = decode

    @property
    def sampler(self):
        return self._index_loader.sampler

    @property
    def dataset(self):
        return self.resident

    def __len__(self) -> int:
        return len(self._index_loader)

    def epoch_indices(self) -> Tuple[torch.Tensor, list]:
        """(flat i64 index tensor on the device, list of batch sizes) for one epoch: one upload.
        Iterating the index loader consumes the global RNG exactly like the reference's loader."""
        batches = [b for b in iter(self._index_loader)]
        if not batches:
            return torch.empty(0, dtype=torch.int64, device=self.resident.device), []
        flat = torch.cat(batches).to(self.resident.device, non_blocking=True)
        return flat, [len(b) for b in batches]

    def epoch_indices_all_ranks(self, epoch: int) -> Optional[torch.Tensor]:
        """i64[world, samples_per_rank] on the device: every rank's shard of `epoch` (DistributedSampler is a pure
        function of (seed, epoch, rank)); None when the sampler is not a DistributedSampler."""
        smp = self._index_loader.sampler
        if not isinstance(smp, DistributedSampler):
            return None
        rows = []
        for r in range(smp.num_replicas):
            sr = DistributedSampler(_IndexOnly(len(self.resident)), num_replicas=smp.num_replicas, rank=r,
                                    shuffle=smp.shuffle, seed=smp.seed, drop_last=smp.drop_last)
            sr.set_epoch(epoch)
            rows.append(torch.tensor(list(iter(sr)), dtype=torch.int64))
        return torch.stack(rows).to(self.resident.device)

    def index_batches(self) -> Iterator[torch.Tensor]:
        """The epoch's batches as int64 index tensors on the device."""
        flat, sizes = self.epoch_indices()
        return iter(torch.split(flat, sizes)) if sizes else iter(())

    def __iter__(self):
        r = self.resident
        for idx in self.index_batches():
            x, y = ops.gather_decode(r.images, r.labels, idx)
            yield x.view(-1, 1, 28, 28), y


def create_data_loaders(config, comm=None, dataset_loc: str = '.', device=None, group=None
                        ) -> Tuple[DeviceBatchLoader, DeviceBatchLoader]:
    """Mirror of the reference function (mp:370-409): same datasets, same sampler arguments
    (`seed=42`, `shuffle=True`), train loader `shuffle=False` + sampler, test loader
    `shuffle=True` over the whole test set on every rank."""
    batch_size = config["trainer"]["batch_size"]
    rank = config["trainer"]["rank"]
    world_size = config["trainer"]["world_size"]
    transform = True                                  # the fused decode == ToTensor + Normalize
    train_dataset = MNISTNetCDF(root_dir=dataset_loc, is_train=True, transforms=transform, comm=comm,
                                device=device, rank=rank, world_size=world_size, group=group)
    test_dataset = MNISTNetCDF(root_dir=dataset_loc, is_train=False, transforms=transform, comm=comm,
                               device=device, rank=rank, world_size=world_size, group=group)
    sampler = DistributedSampler(_IndexOnly(len(train_dataset)), num_replicas=world_size, rank=rank,
                                 shuffle=True, seed=42)
    train_loader = DeviceBatchLoader(train_dataset, batch_size=batch_size, shuffle=False, sampler=sampler)
    test_loader = DeviceBatchLoader(test_dataset, batch_size=batch_size, shuffle=True)
    return train_loader, test_loader
