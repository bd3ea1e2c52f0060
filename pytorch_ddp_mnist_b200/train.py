"""`mnist_pnetcdf_cpu_mp.py`'s script surface on the B200 engine: `configure()`, `init_parallel()`,
`create_data_loaders()`, `create_model()`, `main()`, the same CLI flags, the same printed lines and
the same `model.pt`.

What differs from the reference's `main()` (mp:425-486), by design:
  * the device is the rank's GPU (the reference pins `cpu`, mp:435);
  * the step runs through `TrainEngine` (fused kernels, one flat gradient bucket, one allreduce);
  * the loss is NOT read back every step (`batch_loss.item()`, mp:464, would serialise the device):
    `epoch_loss += batch_loss / B` is accumulated on the device and read once per epoch.  With
    `--progress N` the tqdm description is refreshed every N steps (one sync each time).
  * `finalize()` is called at exit (the reference never destroys its process group).
"""
from __future__ import annotations

import argparse
import os
from typing import Callable, Optional, Tuple

import torch
import torch.distributed as dist
from torch import nn

from . import wireup
from .dataset import DeviceBatchLoader, create_data_loaders
from .engine import PARAM_KEYS, TrainEngine
from .model import B200MLP, create_model

__all__ = ["configure", "init_parallel", "create_data_loaders", "create_model", "main", "run"]


def configure(argv=None):
    """Same flags and the same nested dict as the reference (mp:274-309).  `--data_path` is live
    here (directory holding the .nc files; the reference parses it and then hard-codes '.')."""
    parser = argparse.ArgumentParser(description="Evaluate cost of reading input files")
    add_arg = parser.add_argument
    add_arg("--wireup_method", type=str, default="nccl-slurm", choices=list(wireup.METHODS),
            help="Choose backend for distributed environment initialization")
    add_arg("--data_path", type=str, default=None, help="File path to training samples")
    add_arg("--data_limit", type=int, default=None, help="Max number of samples to be used")
    add_arg("--batch_size", type=int, default=None, help="Batch size")
    add_arg("--n_epochs", type=int, default=None, help="Number of epochs")
    add_arg("--num_workers", type=int, default=None, help="Number of subprocesses to use for data loading")
    add_arg("--parallel", action='store_true', help="To run in parallel")
    add_arg("--hdf5", action='store_true', help="Read from HDF5 files")
    add_arg("--local_rank", "--local-rank", type=int, default=None,
            help="accepted in both spellings (torch.distributed.launch passes --local-rank; tg:123)")
    add_arg("--progress", type=int, default=0, help="refresh the progress line every N steps (0 = off)")
    add_arg("--seed", type=int, default=None, help="torch.manual_seed before create_model() (reference: unseeded)")
    add_arg("--model", type=str, default="ref_mlp", choices=["ref_mlp", "json_convnet"],
            help="ref_mlp = the reference's create_model() (default); json_convnet = the ConvNet BASELINE.json names "
                 "(conv 1->32, conv 32->64, maxpool, fc 9216->128, fc 128->10; Adadelta lr 1.0, StepLR gamma 0.7) - "
                 "NOT in the reference")
    add_arg("--gamma", type=float, default=0.7, help="json_convnet only: StepLR(step_size=1) decay per epoch")
    args = parser.parse_args(argv)

    config = {"trainer": {}, "data": {}}
    config["trainer"]["batch_size"] = 128
    config["trainer"]["wireup_method"] = args.wireup_method
    config["trainer"]["parallel"] = args.parallel
    config["trainer"]["device"] = 0
    config["trainer"]["n_epochs"] = 1
    config["trainer"]["num_workers"] = 0
    config["trainer"]["progress"] = args.progress
    config["trainer"]["seed"] = args.seed
    config["trainer"]["model"] = args.model
    config["trainer"]["gamma"] = args.gamma
    config["data"]["limit"] = None
    config["data"]["label_map"] = [0, 1, 0, 0, 2, 3, 1, 4]
    config["data"]["hdf5"] = args.hdf5
    config["data"]["path"] = args.data_path if args.data_path is not None else "."
    if args.data_limit is not None: config["data"]["limit"] = args.data_limit
    if args.batch_size is not None: config["trainer"]["batch_size"] = args.batch_size
    if args.n_epochs is not None: config["trainer"]["n_epochs"] = args.n_epochs
    if args.num_workers is not None: config["trainer"]["num_workers"] = args.num_workers
    if args.local_rank is not None: os.environ.setdefault("LOCAL_RANK", str(args.local_rank))
    return config


def init_parallel(config):
    """mp:311-367.  No CUDA -> this library cannot run (the reference falls back to gloo+cpu here)."""
    ngpu_per_node = torch.cuda.device_count()
    if not torch.cuda.is_available():
        raise RuntimeError("pytorch_ddp_mnist_b200 needs a CUDA device: there is no CPU fallback")
    rank, world_size = 0, 1
    if config["trainer"]["parallel"]:
        method = config["trainer"]["wireup_method"]
        # `gloo`: the reference's gloo env discovery (OMPI_COMM_WORLD_* / PMIx / PMI, mp:213-254) is kept; only the
        # transport becomes NCCL, because the data is on the GPUs
        comm = wireup.distributed(method, backend="nccl" if method == "gloo" else None)
        rank, world_size, local_rank = comm.get_rank(), comm.get_size(), comm.get_local_rank()
    else:
        comm, local_rank = None, config["trainer"]["device"]
    device = torch.device("cuda:" + str(local_rank))
    torch.cuda.set_device(device)
    config["trainer"]["device"] = device
    config["trainer"]["rank"] = rank
    config["trainer"]["world_size"] = world_size
    if rank == 0:
        import socket
        print('------------------------------------------------------------------')
        print('\n======== MNIST data-parallel training, B200 engine (%s) ========' %
              ('parallel' if config["trainer"]["parallel"] else 'serial'))
        print("%-32s: %s" % ('Host name', socket.gethostname()))
        print("%-32s: %d" % ('Number of processes', world_size))
        print("%-32s: %d" % ('number of GPUs per node', ngpu_per_node))
        print("%-32s: %s" % ('Rank 0 GPU device', device))
        print("%-32s: %s" % ('Input file format', 'netCDF CDF-5, resident in HBM'))
        print("%-32s: %d" % ('Number of epochs', config["trainer"]["n_epochs"]))
        print('------------------------------------------------------------------')
    return comm


def main(model: nn.Module, train_loader: DeviceBatchLoader, test_loader: DeviceBatchLoader, config,
         mask_fn: Optional[Callable[[int, int], torch.Tensor]] = None) -> nn.Module:
    """The reference's `main()` (mp:425-486) on the engine.  Returns the model with the trained
    parameters (no `module.` prefix, like the reference's `model.module`).
    `mask_fn(train_step_number, batch) -> {0,1}[batch,128]` injects dropout masks (parity runs)."""
    rank = config["trainer"]["rank"]
    world_size = config["trainer"].get("world_size", 1)
    device = config["trainer"]["device"]
    if not isinstance(device, torch.device):
        device = torch.device(device)
    epochs = config["trainer"]["n_epochs"]
    progress = config["trainer"].get("progress", 0)
    batch = config["trainer"]["batch_size"]

    model = model.to(device)
    eng = TrainEngine(device, batch, lr=0.01, world_size=world_size, rank=rank)
    eng.load_state_dict(model.state_dict())
    eng.broadcast_params(0)                              # DistributedDataParallel(model), mp:439
    if world_size > 1 and batch <= 512:
        eng.enable_fused_ddp(strict=False)               # gradient average inside the step kernel (NVLink peer memory)
                                                         # when the job is one host, <= 8 ranks, all-pairs P2P; else NCCL
    train_res, test_res = train_loader.resident, test_loader.resident
    step_no = 0
    for i in range(epochs):
        model.train()
        train_loader.sampler.set_epoch(i)                # mp:449
        eng.zero_epoch_loss()
        flat, sizes = train_loader.epoch_indices()
        all_ranks = train_loader.epoch_indices_all_ranks(i) if eng.comm is not None else None
        # consecutive equal-sized batches go to the device as ONE launch of the persistent kernel
        # (`progress` splits the epoch into chunks so a loss can be shown); the short last batch is its own launch
        pos, n = 0, 0
        while n < len(sizes):
            b = sizes[n]
            m = n
            while m < len(sizes) and sizes[m] == b and (not progress or m - n < progress):
                m += 1
            k = m - n
            masks = None
            if mask_fn is not None:
                masks = torch.stack([mask_fn(step_no + t, b) for t in range(k)]).to(device, torch.uint8).contiguous()
            ia = all_ranks[:, pos:pos + k * b].contiguous() if all_ranks is not None else None
            eng.train_steps(train_res, flat[pos:pos + k * b], b, k, masks=masks, idx_all=ia)
            step_no += k
            pos += k * b
            n = m
            if progress and rank == 0:
                print(f'training batch_loss={eng.last_batch_loss():.4f} [{n}/{len(sizes)}]')
        eng.check_comm()
        epoch_loss = eng.epoch_loss()
        with torch.no_grad():
            model.eval()
            eng.zero_epoch_loss()
            for idx in test_loader.index_batches():      # whole test set on every rank, mp:402-407
                eng.eval_step(test_res, idx)
            val_loss = eng.epoch_loss()
        print(f"Epoch={i}, train_loss={epoch_loss:.4f}, val_loss={val_loss:.4f}")
    sd = eng.state_dict()
    with torch.no_grad():
        for k, v in model.state_dict().items():
            v.copy_(sd[k])
    return model


def main_convnet(train_loader: DeviceBatchLoader, test_loader: DeviceBatchLoader, config):
    """The recipe BASELINE.json describes for its ConvNet (NOT in the reference: there is nothing to cite; oracle =
    oracle/convnet_oracle.py, parity unpinned): per epoch one pass over this rank's shard with Adadelta(lr=1.0), the
    bucket all-reduced over NCCL at ws > 1, then the whole test set (loss + accuracy) on every rank and
    `StepLR(step_size=1, gamma)`.  Same loaders, sampler and printed line shape as `main()`; returns the engine."""
    from .convnet import ConvNetEngine, init_params_like_torch
    rank = config["trainer"]["rank"]
    world_size = config["trainer"].get("world_size", 1)
    device = torch.device(config["trainer"]["device"])
    batch = config["trainer"]["batch_size"]
    eng = ConvNetEngine(device, batch, lr=1.0, world_size=world_size, rank=rank, tf32=True)
    eng.load_state_dict(init_params_like_torch())
    eng.broadcast_params(0)
    train_res, test_res = train_loader.resident, test_loader.resident
    for i in range(config["trainer"]["n_epochs"]):
        train_loader.sampler.set_epoch(i)
        eng.loss.zero_()
        for idx in train_loader.index_batches():
            eng.train_step(train_res, idx)
        epoch_loss = float(eng.loss[1].item())
        eng.loss.zero_()
        correct, total = 0.0, 0
        for idx in test_loader.index_batches():
            correct += eng.accuracy(test_res, idx) * idx.numel()     # eval forward; accumulates loss[1] like mp:481
            total += idx.numel()
        val_loss = float(eng.loss[1].item())
        print(f"Epoch={i}, train_loss={epoch_loss:.4f}, val_loss={val_loss:.4f}, val_acc={correct / max(total, 1):.4f}")
        eng.scale_lr(config["trainer"].get("gamma", 0.7))
    return eng


def run(argv=None) -> None:
    """The reference's `__main__` block (mp:489-511)."""
    config = configure(argv)
    comm = init_parallel(config)
    rank = config["trainer"]["rank"]
    if config["trainer"]["seed"] is not None:
        torch.manual_seed(config["trainer"]["seed"])
    train_loader, test_loader = create_data_loaders(config, None, dataset_loc=config["data"]["path"],
                                                    device=config["trainer"]["device"])
    if config["trainer"].get("model") == "json_convnet":
        eng = main_convnet(train_loader, test_loader, config)
        if rank == 0:
            torch.save(eng.state_dict(), 'convnet.pt')
        if comm is not None:
            comm.finalize()
        return
    model = main(model=create_model(), train_loader=train_loader, test_loader=test_loader, config=config)
    if rank == 0:
        torch.save({k: v.detach().cpu() for k, v in model.state_dict().items()}, 'model.pt')
    if comm is not None:
        comm.finalize()
