"""`bench.py --config scaled6m`: BASELINE.json configs[4] - the scaled synthetic CDF-5 file (6,000,000 x 28 x 28 uint8,
4.7 GB) loaded into HBM and trained for ONE FULL EPOCH at batch 64 per rank, next to the reference's per-sample
loader (`mnist_pnetcdf_cpu_mp.py:39-49`, driven by `:396-401` and the loop at `:453-466`) timed on the same file.

What is measured (one JSON line, same contract as bench.py):
  loader     `ResidentDataset.from_file`: every rank `pread`s a contiguous 1/N slab through two pinned staging buffers
             (read of chunk i+1 overlaps the H2D copy of chunk i), then one NCCL all-gather replicates the u8 dataset
             (a rank's shard `perm[rank::N]` is scattered over the whole file).  Reported: seconds, file->HBM GB/s
             (slab bytes / read+copy time), all-gather GB/s (bus), and the same load with the page cache warm.
  value      whole-job images/s of the epoch (N x ceil(6M/N/64) steps incl. the short last batch), dataset resident,
             ONE launch of the persistent kernel per rank for all full batches + one for the last batch; CUDA events,
             max over ranks.  The loader time is reported beside it, and `epoch_with_load_s` adds the two.
  e2e        the same epoch's batches fed from pinned HOST memory (`TrainEngine.train_steps_host`), capped at
             `--steps` batches per rank when that is smaller than the epoch (pinning 4.7 GB on one rank is the cap's
             reason); says how many steps it ran.
  cpu_baseline   oracle/ref_port.py (the reference's loop incl. its per-sample netCDF reads) on the SAME file, bounded
             sample of steps.
`--samples` shrinks the file (default 6,000,000) for quick runs.
"""
from __future__ import annotations

import json
import os
import sys
import tempfile
import threading
import time


def _part_worker(path, n, seed, part, nparts):
    from pytorch_ddp_mnist_b200 import synth
    synth.write_synth_file_part(path, n, seed, part, nparts, False)


def _make_file(path: str, n: int, seed: int, part: int = 0, nparts: int = 1, procs: int = 0) -> float:
    """This process's share (part of nparts) of the file, written by `procs` worker processes (chunks are independent).
    The header must exist already (write_synth_file_part(create=True))."""
    import multiprocessing as mp
    t0 = time.perf_counter()
    procs = procs or max(1, min(16, (os.cpu_count() or 8) // max(nparts, 1)))
    ctx = mp.get_context("fork")
    ws = [ctx.Process(target=_part_worker, args=(path, n, seed, part * procs + k, nparts * procs)) for k in range(procs)]
    for w in ws:
        w.start()
    for w in ws:
        w.join()
        if w.exitcode != 0:
            raise RuntimeError("synthetic file writer failed")
    return time.perf_counter() - t0


def _timed_load(path, dev, rank, world, dist, torch):
    """One `ResidentDataset.from_file` with its phases timed on the host clock (it is a host-driven pipeline)."""
    from pytorch_ddp_mnist_b200.dataset import ResidentDataset
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    stats = {}
    t0 = time.perf_counter()
    res = ResidentDataset.from_file(path, dev, rank=rank, world_size=world, stats=stats)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    out = {"seconds": dt, **stats}
    if dist is not None:                                      # slowest rank
        t = torch.tensor([dt, stats.get("read_copy_s", 0.0), stats.get("allgather_s", 0.0)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out["seconds"], out["read_copy_s"], out["allgather_s"] = [float(v) for v in t.tolist()]
    return res, out


def run(args, clock_sampler_cls, metric) -> None:
    import numpy as np
    import torch
    import torch.distributed as dist
    from torch.utils.data import DistributedSampler
    from pytorch_ddp_mnist_b200 import engine as E
    from pytorch_ddp_mnist_b200.dataset import _IndexOnly

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    n = int(os.environ.get("DDPM_SCALED_SAMPLES", "6000000"))
    B = args.batch
    if args.impl == "reference":
        if rank == 0:
            _reference_arm(args, n, metric)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dd = None
    if world > 1:
        dist.init_process_group("nccl", init_method="env://", device_id=dev)
        dd = dist

    # ---- the file (synthetic generator of SURVEY.md §8(d), seed 1234, CDF-5): rank 0 lays it out, then every rank
    #      writes its share of the 100,000-sample chunks with a few worker processes (chunks have independent generators)
    box = [None, False]
    if rank == 0:
        d = os.environ.get("DDPM_SCALED_DIR") or tempfile.mkdtemp(prefix="scaled6m_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
        path = os.path.join(d, "mnist_train_images.nc")
        fresh = not os.path.exists(path)
        if fresh:
            from pytorch_ddp_mnist_b200 import synth
            synth.write_synth_file_part(path, n, 1234, 0, 1, True)
        box = [path, fresh]
    if world > 1:
        dist.broadcast_object_list(box, src=0)
    path, fresh = box
    gen_s = 0.0
    if fresh:
        gen_s = _make_file(path, n, 1234, rank, world)
        if world > 1:
            dist.barrier()
    file_bytes = os.path.getsize(path)

    # ---- loader: first load (page cache as the writer left it), second load (warm)
    res, load1 = _timed_load(path, dev, rank, world, dd, torch)
    del res
    torch.cuda.empty_cache()
    res, load2 = _timed_load(path, dev, rank, world, dd, torch)
    slab_bytes = -(-n // world) * 785
    for ld in (load1, load2):
        ld["file_to_hbm_gbs_per_rank"] = slab_bytes / max(ld.get("read_copy_s", ld["seconds"]), 1e-9) / 1e9
        ld["file_to_hbm_gbs_job"] = ld["file_to_hbm_gbs_per_rank"] * world
        if world > 1 and ld.get("allgather_s"):
            ld["allgather_bus_gbs"] = n * 785 * (world - 1) / world / ld["allgather_s"] / 1e9

    # ---- one full epoch: the reference's epoch 0 (DistributedSampler(seed=42).set_epoch(0), mp:390-394, 449)
    sampler = DistributedSampler(_IndexOnly(n), num_replicas=world, rank=rank, shuffle=True, seed=42)
    sampler.set_epoch(0)
    shard = torch.tensor(list(iter(sampler)), dtype=torch.int64)
    per_rank = shard.numel()
    full, tail = per_rank // B, per_rank % B
    shard_dev = shard.to(dev)
    idx_all = None
    torch.manual_seed(0)
    eng = E.TrainEngine(dev, B, lr=0.01, world_size=world, rank=rank)
    eng.load_state_dict(E.init_params_like_reference())
    eng.broadcast_params()
    if world > 1:
        eng.enable_fused_ddp()
        if world <= 4:                                         # dz1 exchange needs every rank's batches
            rows = []
            for r in range(world):
                sr = DistributedSampler(_IndexOnly(n), num_replicas=world, rank=r, shuffle=True, seed=42)
                sr.set_epoch(0)
                rows.append(torch.tensor(list(iter(sr)), dtype=torch.int64))
            idx_all = torch.stack(rows).to(dev)

    def epoch():
        ia = idx_all[:, :full * B].contiguous() if idx_all is not None else None
        eng.train_steps(res, shard_dev[:full * B], B, full, idx_all=ia)
        if tail:
            ib = idx_all[:, full * B:].contiguous() if idx_all is not None else None
            eng.train_steps(res, shard_dev[full * B:], tail, 1, idx_all=ib)
        return 2 if tail else 1

    eng.train_steps(res, shard_dev[:50 * B], B, 50, idx_all=idx_all[:, :50 * B].contiguous() if idx_all is not None else None)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    smp = clock_sampler_cls(local_rank)
    smp.start()
    smp.mark()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    launches = epoch()
    e1.record()
    torch.cuda.synchronize()
    smp.mark()
    clocks = smp.stop()
    eng.check_comm()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    steps = full + (1 if tail else 0)
    images = per_rank * world
    value = images / (ms * 1e-3)
    final_loss = eng.last_batch_loss()

    # ---- the same epoch's batches from pinned host memory (capped)
    hsteps = min(full, max(args.steps, 1000))
    rows = shard[:hsteps * B].numpy()
    from pytorch_ddp_mnist_b200 import cdf5
    hdr = cdf5.read_header(path)
    mm = np.memmap(path, dtype=np.uint8, mode="r")
    iv, lv = hdr.vars["images"], hdr.vars["labels"]
    imgs = mm[iv.begin:iv.begin + n * 784].reshape(n, 784)
    labs = mm[lv.begin:lv.begin + n]
    t0 = time.perf_counter()
    hx = torch.from_numpy(np.ascontiguousarray(imgs[rows])).view(hsteps, B, 784).pin_memory()
    hy = torch.from_numpy(np.ascontiguousarray(labs[rows])).view(hsteps, B).pin_memory()
    gather_s = time.perf_counter() - t0
    hloss = torch.full((hsteps,), float("nan")).pin_memory()
    eng.train_steps_host(hx[:20], hy[:20], B, 20, host_loss_steps=hloss[:20])
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    e0.record()
    eng.train_steps_host(hx, hy, B, hsteps, host_loss_steps=hloss)
    e1.record()
    torch.cuda.synchronize()
    hms = max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3)
    eng.check_feed()
    if world > 1:
        t = torch.tensor([hms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        hms = float(t.item())
    assert bool(torch.isfinite(hloss).all())
    e2e = {"value": hsteps * B * world / (hms * 1e-3), "unit": "images/s", "h2d_bytes_per_step": B * 785 + 8,
           "d2h_bytes_per_step": 4, "steps": hsteps, "api": "TrainEngine.train_steps_host (ddpm_train_steps_host_fed)",
           "host_gather_s": gather_s,
           "note": f"first {hsteps} of the epoch's {full} full batches per rank, fed from pinned host memory"}

    P = 473_088
    bytes_step = B * (784 + 8 + 1) + 2 * P
    step_s = ms * 1e-3 / steps
    peaks = os.path.join(os.path.dirname(os.path.abspath(__file__)), "MEASURED_PEAKS.json")
    hbm = float(json.load(open(peaks))["hbm_gbs"]) if os.path.exists(peaks) else 6650.0
    line = {
        "metric": metric, "value": value, "unit": "images/s", "n_gpus": world, "steps": steps, "warmup": 50,
        "ms_per_step": ms / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "model": "ref_mlp",
        "config": {"workload": f"scaled6m: BASELINE.json configs[4] - {n}-sample synthetic CDF-5 file ({file_bytes / 1e9:.2f} GB), "
                               f"one full epoch at batch {B}/rank ({steps} steps/rank, last batch {tail or B}), ref_mlp",
                   "dataset_samples": n, "file_bytes": file_bytes, "file_generation_s": gen_s, "batch_per_rank": B,
                   "final_batch_loss": final_loss,
                   "l2": "4.7 GB resident dataset, every row read once per epoch"},
        "epoch_s": ms * 1e-3, "epoch_with_load_s": ms * 1e-3 + load2["seconds"],
        "loader": {"first": load1, "warm": load2,
                   "path": "pread slab -> 2 pinned staging buffers -> cudaMemcpyAsync (double-buffered) -> NCCL all-gather"},
        "gpu_launches": launches, "clocks": clocks, "e2e": e2e,
        "roofline": {"bound": "hbm", "achieved": bytes_step / step_s / 1e9, "peak": hbm, "unit": "GB/s",
                     "frac": bytes_step / step_s / 1e9 / hbm, "traffic": None, "alg_bytes_per_step": bytes_step,
                     "kernel": "train_persist_kernel, per step of the epoch-long launch"},
    }
    if rank == 0:
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = _cpu_sample(os.path.dirname(path), B, min(args.cpu_baseline_steps, 150), 2)
        print(json.dumps(line), flush=True)
        if not os.environ.get("DDPM_SCALED_DIR") and not os.environ.get("DDPM_SCALED_KEEP"):
            try:
                os.remove(path)
                os.rmdir(os.path.dirname(path))
            except OSError:
                pass
    if world > 1:
        sys.stdout.flush()
        threading.Timer(20.0, lambda: os._exit(0)).start()
        torch.cuda.synchronize()
        dist.barrier()
        os._exit(0)


def _cpu_sample(root, batch, steps, ws):
    """The reference's loop (port) reading THIS file sample by sample; needs the test file's name to exist too."""
    from oracle import ref_port
    from pytorch_ddp_mnist_b200 import synth
    test = os.path.join(root, "mnist_test_images.nc")
    made = False
    if not os.path.exists(test):
        synth.write_synth_file(test, 1000, 4321)
        made = True
    try:
        r = ref_port.time_port_gloo(root, batch, steps, 5, world_size=ws, timeout_s=420.0)
    finally:
        if made:
            os.remove(test)
    return {"value": r["images_per_s"], "unit": "images/s", "cores": r["cores"], "kind": "port",
            "sample": f"{steps} steps x batch {batch} x {ws} gloo ranks ({r['threads_per_rank']} threads each), per-sample "
                      f"CDF-5 reads from the same file, loader share {r['loader_fraction']:.2f}, {r['seconds']:.1f} s"}


def _reference_arm(args, n, metric):
    d = os.environ.get("DDPM_SCALED_DIR") or tempfile.mkdtemp(prefix="scaled6m_ref_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    path = os.path.join(d, "mnist_train_images.nc")
    if not os.path.exists(path):
        from pytorch_ddp_mnist_b200 import synth
        synth.write_synth_file_part(path, n, 1234, 0, 1, True)
        _make_file(path, n, 1234)
    ws = max(2, args.gpus)
    steps = max(1, min(args.steps, 300))
    r = _cpu_sample(d, args.batch, steps, ws)
    line = {"impl": "reference", "metric": metric, "value": r["value"], "unit": "images/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": 5, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "model": "ref_mlp",
            "config": {"workload": f"scaled6m: {n}-sample synthetic CDF-5 file, reference loop (port), bounded sample",
                       "reference_world_size": ws, "batch_per_rank": args.batch},
            "cpu_baseline": r, "gpu_launches": 0,
            "e2e": {"value": r["value"], "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)
    if not os.environ.get("DDPM_SCALED_DIR"):
        try:
            os.remove(path)
            os.rmdir(d)
        except OSError:
            pass
