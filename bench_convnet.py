"""`bench.py --model json_convnet`: the BASELINE.json "ConvNet" step.

NOT the reference's model: /root/reference defines only the MLP that the default
`bench.py` run measures.  This arm exists because BASELINE.json's configs name the ConvNet; every
number it prints is tagged `json_convnet` and its CPU baseline is the restatement oracle
(oracle/convnet_oracle.py, parity unpinned), not reference code.

Same JSON contract as bench.py: one step = forward + backward + Adadelta on one batch of synthetic
images per rank, dataset resident in HBM for `value`; `e2e` feeds every step's batch from pinned host
memory and reads the loss back; weak scaling with an NCCL all-reduce of the 4.8 MB bucket for N > 1.
"""
from __future__ import annotations

import json
import os
import sys
import threading
import time


def _tf32_peak(torch, dev):
    """Dense TF32 matmul throughput of this GPU (cuBLAS, 8192^3, best of 10): the tensor-pipe denominator
    SURVEY.md §6 asks for (MEASURED_PEAKS.json holds bf16 only)."""
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        a = torch.randn(8192, 8192, device=dev)
        b = torch.randn(8192, 8192, device=dev)
        for _ in range(2):
            a @ b
        torch.cuda.synchronize()
        best = 1e9
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for _ in range(10):
            e0.record()
            a @ b
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        return 2 * 8192 ** 3 / (best * 1e-3) / 1e12
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev


def _cpu_baseline(batch, steps):
    """oracle/convnet_oracle.py (torch CPU fp32, all host threads) on a bounded sample of the same workload."""
    import numpy as np
    import torch
    from oracle import convnet_oracle as CO
    from oracle import mnist_oracle as O
    from pytorch_ddp_mnist_b200 import synth
    im, lb = synth.synth(4096, 7)
    p = CO.flatten_params(CO.init_params(0))
    sq, acc = np.zeros_like(p), np.zeros_like(p)
    rng = np.random.default_rng(0)
    t0 = None
    for t in range(steps + 2):
        if t == 2:
            t0 = time.perf_counter()
        idx = rng.integers(0, 4096, batch).astype(np.int64)
        x, y = O.gather_decode(im, lb, idx)
        m1 = (rng.random((batch, 9216)) < 0.75).astype(np.uint8)
        m2 = (rng.random((batch, 128)) < 0.5).astype(np.uint8)
        _, _, g, _ = CO.forward_backward(CO.unflatten_params(p), x.reshape(batch, 784), y, m1, m2)
        p, sq, acc = CO.adadelta_step(p, CO.flatten_params(g), sq, acc)
    dt = time.perf_counter() - t0
    return {"value": steps * batch / dt, "unit": "images/s", "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"{steps} steps x batch {batch}, torch CPU fp32 restatement of BASELINE.json's layer list "
                      f"(oracle/convnet_oracle.py; the model is not in the reference), {dt:.1f} s"}


FLOP_FWD, FLOP_BWD = 23_984_896, 47_580_416          # per image, SURVEY.md §8(a)
CONV2_FLOP = 21_233_664                              # conv2 forward = conv2 dgrad = conv2 wgrad (dense), per image


def quick_record(torch, dev, B, world, rank, dist, steps=150, warmup=10):
    """Short resident-data measurement of the json_convnet step for the `context.json_convnet` sub-record of the default
    bench line (so that the driver's BENCH/SCALE records carry a ConvNet number at all).  TF32 tensor-core path,
    weak scaling with the NCCL bucket all-reduce at ws > 1; CUDA events, max over ranks."""
    import numpy as np
    from pytorch_ddp_mnist_b200 import convnet as cn, synth
    from pytorch_ddp_mnist_b200.dataset import ResidentDataset
    N = 60_000
    im, lb = synth.synth(N, 20260212)
    res = ResidentDataset.from_arrays(im, lb, dev)
    g = torch.Generator().manual_seed(7 + rank)
    idx = torch.randint(0, N, ((steps + warmup) * B,), generator=g, dtype=torch.int64).to(dev)
    eng = cn.ConvNetEngine(dev, B, world_size=world, rank=rank, tf32=True)
    torch.manual_seed(0)
    eng.load_state_dict(cn.init_params_like_torch())
    eng.broadcast_params()
    for t in range(warmup):
        eng.train_step(res, idx[t * B:(t + 1) * B])
    chunk = 10                                                # CUDA graph of 10 unrolled steps, index buffer refilled per replay
    steps = (steps // chunk) * chunk
    buf = torch.empty(chunk * B, dtype=torch.int64, device=dev)
    graph = eng.capture_steps(res, list(torch.split(buf, B)))
    buf.copy_(idx[:chunk * B])
    graph.replay()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for t in range(warmup, warmup + steps, chunk):
        buf.copy_(idx[t * B:(t + chunk) * B])
        graph.replay()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if dist is not None:
        tt = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms = float(tt.item())
    loss = eng.last_batch_loss()
    peak = _tf32_peak(torch, dev)
    step_s = ms * 1e-3 / steps
    tfl = (FLOP_FWD + FLOP_BWD) * B / step_s / 1e12
    return {"model": "json_convnet (BASELINE.json's ConvNet; NOT in the reference, oracle = restatement, parity unpinned)",
            "value": steps * B * world / (ms * 1e-3), "unit": "images/s", "batch_per_rank": B, "steps": steps,
            "us_per_step": 1e6 * step_s, "dtype": "tf32", "final_batch_loss": loss,
            "roofline": {"bound": "tensor", "achieved": tfl, "peak": peak, "unit": "TFLOP/s", "frac": tfl / peak,
                         "peak_source": "cuBLAS TF32 matmul 8192^3 measured here, best of 10"},
            "full_line": "python bench.py --model json_convnet"}


def run(args, clock_sampler_cls, peaks_fn, metric) -> None:
    import numpy as np
    import torch
    import torch.distributed as dist
    from pytorch_ddp_mnist_b200 import convnet as cn, synth
    from pytorch_ddp_mnist_b200.dataset import ResidentDataset, _IndexOnly
    from torch.utils.data import DistributedSampler

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", init_method="env://", device_id=dev)
    B, K, W = args.batch, min(args.steps, 2000), max(3, min(args.warmup, 50))
    tf32 = not args.fp32
    N = 240_000
    im, lb = synth.synth(N, 20260212)
    res = ResidentDataset.from_arrays(im, lb, dev)
    sampler = DistributedSampler(_IndexOnly(N), num_replicas=world, rank=rank, shuffle=True, seed=42)
    sampler.set_epoch(0)
    shard = torch.tensor(list(iter(sampler)), dtype=torch.int64)
    need = (W + K + 110) * B                                # + the per-phase timing section's 100 fresh batches
    shard_dev = shard.repeat(-(-need // shard.numel()))[:need].to(dev)
    eng = cn.ConvNetEngine(dev, B, world_size=world, rank=rank, tf32=tf32)
    torch.manual_seed(0)
    eng.load_state_dict(cn.init_params_like_torch())
    eng.broadcast_params()

    # the step is 13 (TF32) / 10 (fp32) dependent launches: replayed as a CUDA graph of `chunk` unrolled steps whose
    # index buffer is refilled before every replay (`--no-graph`-style eager launches remain for the tail)
    chunk = 10
    gbuf = torch.empty(chunk * B, dtype=torch.int64, device=dev)
    for t in range(2):
        eng.train_step(res, shard_dev[t * B:(t + 1) * B])
    graph = eng.capture_steps(res, list(torch.split(gbuf, B)))

    def steps(n, start):
        t = start
        while t + chunk <= start + n:
            gbuf.copy_(shard_dev[t * B:(t + chunk) * B])
            graph.replay()
            t += chunk
        while t < start + n:
            eng.train_step(res, shard_dev[t * B:(t + 1) * B])
            t += 1

    steps(W, 0)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    smp = clock_sampler_cls(local_rank)
    smp.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    steps(K, W)
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    clocks = smp.stop()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = K * B * world / (ms * 1e-3)
    final_loss = eng.last_batch_loss()

    # ---- per-phase durations (CUDA events on the launching stream, fresh rows every call)
    reps = 100
    idxs = [shard_dev[i * B:(i + 1) * B] for i in range(reps)]
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]

    def avg_ms(fn):
        for i in range(5):
            fn(idxs[i])
        torch.cuda.synchronize()
        ev[0].record()
        for i in range(reps):
            fn(idxs[i])
        ev[1].record()
        torch.cuda.synchronize()
        return ev[0].elapsed_time(ev[1]) / reps

    t_fwd = avg_ms(lambda ix: cn.step_forward(eng.params, eng.ws, images=res.images, labels=res.labels, idx=ix,
                                              philox=(1, 0), prep_backward=True, loss=eng.loss, tf32=tf32))
    t_bwd = avg_ms(lambda ix: cn.step_backward(eng.params, eng.ws, eng.grads, images=res.images, idx=ix, tf32=tf32,
                                               forward_was_tf32=tf32))
    scratch = [torch.zeros_like(eng.params) for _ in range(3)]
    t_opt = avg_ms(lambda ix: cn.adadelta_step(scratch[0], eng.grads, scratch[1], scratch[2]))
    peak_tf32 = _tf32_peak(torch, dev)
    hbm, which = peaks_fn()
    step_ms = ms / K
    flops = (FLOP_FWD + FLOP_BWD) * B
    roof = {"bound": "tensor", "achieved": flops / (step_ms * 1e-3) / 1e12, "peak": peak_tf32, "unit": "TFLOP/s",
            "frac": flops / (step_ms * 1e-3) / 1e12 / peak_tf32, "traffic": None,
            "kernel": "whole step (13 kernels; conv2 forward / data gradient / weight gradient on tcgen05 TF32 are "
                      "89 % of the flops)",
            "peak_source": "measured here: cuBLAS TF32 matmul 8192^3, best of 10 (MEASURED_PEAKS.json has bf16 only)",
            "note": "algorithmic flops = 71,565,312 per image (SURVEY.md §8(a)); at batch 64 the step is 4.6 GFLOP: "
                    "gather/epilogue/launch latency, not the tensor pipe, sets the time"}
    kernels = [{"name": "forward (conv1, conv2+pool on tcgen05, fc1 split-K, head)", "ms": t_fwd,
                "tflops": FLOP_FWD * B / (t_fwd * 1e-3) / 1e12},
               {"name": "backward (fc1, conv2 dgrad + wgrad on tcgen05, conv1, reductions)", "ms": t_bwd,
                "tflops": FLOP_BWD * B / (t_bwd * 1e-3) / 1e12},
               {"name": "adadelta_kernel (7 x 4.8 MB of traffic)", "ms": t_opt, "gbs": 7 * 4 * 1_199_882 / t_opt / 1e6,
                "hbm_peak_gbs": hbm, "hbm_peak_source": which}]

    # ---- e2e: every step's batch comes from pinned host memory, the loss goes back
    esteps = min(K, 500)
    rows = np.resize(shard.numpy(), (esteps + 6) * B) % N
    host_x = torch.from_numpy(im.reshape(-1, 784)[rows]).pin_memory()
    host_y = torch.from_numpy(lb[rows]).pin_memory()
    xd = [torch.empty(B, 784, dtype=torch.uint8, device=dev) for _ in range(2)]
    yd = [torch.empty(B, dtype=torch.uint8, device=dev) for _ in range(2)]
    ident = torch.arange(B, dtype=torch.int64, device=dev)
    loss_host = torch.zeros(esteps + 5, 2).pin_memory()

    class _Staged:
        images, labels = None, None

    # copies ride their own stream one batch ahead of the compute stream (double buffer, event-ordered)
    copy_stream = torch.cuda.Stream(device=dev)
    copied = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]
    main = torch.cuda.current_stream()

    def stage(t):
        b = t & 1
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed[b])                 # the step that last read buffer b has finished
            xd[b].copy_(host_x[t * B:(t + 1) * B], non_blocking=True)
            yd[b].copy_(host_y[t * B:(t + 1) * B], non_blocking=True)
            copied[b].record(copy_stream)

    def one(t):
        b = t & 1
        stage(t + 1)                                            # next batch's H2D overlaps this step
        main.wait_event(copied[b])
        _Staged.images, _Staged.labels = xd[b], yd[b]
        eng.train_step(_Staged, ident)
        consumed[b].record(main)
        loss_host[t].copy_(eng.loss, non_blocking=True)

    for b in range(2):
        consumed[b].record(main)
    stage(0)
    for t in range(5):
        one(t)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ev[0].record()
    for t in range(5, 5 + esteps):
        one(t)
    ev[1].record()
    torch.cuda.synchronize()
    ems = ev[0].elapsed_time(ev[1])
    if world > 1:
        tt = torch.tensor([ems], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ems = float(tt.item())
    assert bool(torch.isfinite(loss_host[5:5 + esteps, 0]).all())
    e2e = {"value": esteps * B * world / (ems * 1e-3), "unit": "images/s", "h2d_bytes_per_step": B * 785,
           "d2h_bytes_per_step": 8, "steps": esteps,
           "api": "ConvNetEngine.train_step on a staged host batch (2 H2D copies one batch ahead on a copy stream, 14 launches, async D2H of the loss)"}

    line = {"metric": metric, "value": value, "unit": "images/s", "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "tf32" if tf32 else "f32", "data": "synthetic",
            "config": {"workload": f"json_convnet (BASELINE.json's ConvNet: conv 1->32, conv 32->64, maxpool, fc 9216->128, "
                                   f"fc 128->10, nll_loss, Adadelta) training step, batch {B}/rank - NOT in the reference "
                                   f"(SURVEY.md §0); oracle = restatement, parity unpinned",
                       "batch_per_rank": B, "dataset_samples": N, "final_batch_loss": final_loss,
                       "l2": "resident u8 dataset 188 MB (> 126 MB L2), fresh rows every step",
                       "gradient_exchange": "NCCL all-reduce of the 4,799,528-byte bucket between backward and Adadelta (ws > 1)",
                       "step_launch": "CUDA graph of 10 unrolled steps, index buffer refilled per replay"},
            "gpu_launches": K * (13 if tf32 else 10), "clocks": clocks, "e2e": e2e, "roofline": roof, "kernels": kernels}
    if rank == 0:
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = _cpu_baseline(B, 150)
        print(json.dumps(line), flush=True)
    if world > 1:
        sys.stdout.flush()
        threading.Timer(20.0, lambda: os._exit(0)).start()
        torch.cuda.synchronize()
        dist.barrier()
        os._exit(0)
