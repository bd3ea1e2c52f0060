#!/usr/bin/env python
"""Benchmark of the hot path: one data-parallel MNIST training step (reference model `ref_mlp`,
`mnist_pnetcdf_cpu_mp.py:453-466`) on synthetic MNIST-shaped data.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch 64] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

Prints ONE JSON line on rank 0.  See DESIGN.md §6 for what every field means.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
if REPO not in sys.path:
    sys.path.insert(0, REPO)

METRIC = "MNIST-ConvNet train images/sec at 1/2/4/8 B200 + % roofline vs CPU gloo ref"   # BASELINE.json
UNIT = "images/s"
BENCH_N = 240_000          # 188 MB of u8 pixels: larger than the 126 MB L2
BENCH_SEED = 1234
CHUNK = 100                # steps unrolled per CUDA graph


def _ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum of one 1-step launch of train_persist_kernel, from the
    committed `ncu --set full` capture (profiles/r01_v2_ncu_full_summary.csv); None if absent."""
    import csv
    p = os.path.join(REPO, "profiles", "r01_v2_ncu_full_summary.csv")
    try:
        rows = list(csv.reader(open(p)))
        hdr, units, r = rows[0], rows[1], rows[2]
        mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        tot = 0.0
        for name in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            i = hdr.index(name)
            tot += float(r[i]) * mult[units[i]]
        return tot
    except Exception:
        return None


def _peaks():
    p = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ------------------------------------------------------------------------------------ clocks
class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons of one GPU through NVML while the timed region runs."""

    def __init__(self, index: int, period_s: float = 0.02):
        super().__init__(daemon=True)
        self.index, self.period = index, period_s
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop_evt = threading.Event()
        self.err = None

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {
                nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
                nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: "hw_power_brake",
            }
            while not self._stop_evt.is_set():
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
                time.sleep(self.period)
        except Exception as e:                                   # NVML missing: report it
            self.err = repr(e)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s), **({"error": self.err} if self.err else {})}


# ------------------------------------------------------------------------------------ reference arm
def run_reference_arm(args) -> None:
    """`--impl reference`: the reference's own CPU path (gloo DDP, world_size 2, per-sample netCDF
    loader) on the host cores, via oracle/ref_port.py (the Python reference cannot travel to the
    GPU box; the port is checked against it in tests/test_ref_port.py)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import ref_port
    from pytorch_ddp_mnist_b200 import synth
    steps = max(1, min(args.steps, 400))            # bounded sample: ~10 ms/step on CPU
    warmup = max(1, min(args.warmup, 20))
    with tempfile.TemporaryDirectory() as d:
        synth.write_default_files(d, 60_000, 1_000)
        r = ref_port.time_port_gloo(d, args.batch, steps, warmup, world_size=2, port=29700 + os.getpid() % 200)
    v = r["images_per_s"]
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": 1e3 * r["seconds"] / steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": _config(args, extra={"reference_world_size": 2}),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": r["cores"], "kind": "port",
                         "sample": f"{steps} steps x batch {args.batch} x 2 gloo ranks "
                                   f"({r['threads_per_rank']} threads each) of epoch 0, synthetic 60000-sample CDF-5 file; "
                                   f"loader share {r['loader_fraction']:.2f}"},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def _config(args, extra=None):
    c = {"workload": f"ref_mlp (784-128-128-10 MLP, the model the reference actually defines) training step, "
                     f"batch {args.batch}/rank, CrossEntropy + SGD(lr=0.01), dropout 0.2; BASELINE.json configs[1] "
                     f"(the 'ConvNet' it names does not exist in the reference, SURVEY.md §0)",
         "batch_per_rank": args.batch, "dataset_samples": BENCH_N,
         "l2": "resident u8 dataset is 188 MB (> 126 MB L2) and every step gathers rows not touched before; "
               "weights stay L2-resident as in any training loop",
         "steps_per_launch_or_graph": CHUNK,
         "step_kernel": "three-kernel (v1)" if args.three_kernel or args.batch > 512 else "persistent (v2)",
         "gradient_exchange": "NCCL allreduce between kernels" if args.no_fused else
                              "in-kernel over NVLink peer memory (flag-in-data pushes, sums in fixed rank order; fc1's "
                              "gradient exchange replaced by the dz1 row exchange), ws > 1 only"}
    if args.batch > 512:
        c["tensor_cores"] = "fc1 forward + dW1 on tcgen05 TF32" if args.tf32 else "off (fp32 FFMA)"
    if extra:
        c.update(extra)
    return c


# ------------------------------------------------------------------------------------ our arm
def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20000)
    ap.add_argument("--warmup", type=int, default=200)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-fused", action="store_true", help="ws > 1: NCCL allreduce between kernels instead of the fused peer-memory exchange")
    ap.add_argument("--three-kernel", action="store_true", help="use the v1 three-kernel step instead of the persistent kernel")
    ap.add_argument("--tf32", action="store_true", help="batch > 512 only: fc1 forward and dW1 on tcgen05 TF32 tensor cores (fp32 accumulate)")
    ap.add_argument("--cpu-baseline-steps", type=int, default=150)
    ap.add_argument("--model", default="ref_mlp", choices=["ref_mlp", "json_convnet"],
                    help="ref_mlp = the model the reference defines (default, the graded path); json_convnet = the ConvNet "
                         "BASELINE.json names, which is not in the reference (bench_convnet.py)")
    ap.add_argument("--fp32", action="store_true", help="json_convnet only: fp32 FFMA kernels instead of tcgen05 TF32")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.model == "json_convnet":
        import bench_convnet
        if args.impl == "reference":            # no reference code exists for this model: the CPU arm is the restatement
            if int(os.environ.get("RANK", "0")) == 0:
                r = bench_convnet._cpu_baseline(args.batch, 150)
                print(json.dumps({"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT,
                                  "n_gpus": args.gpus, "steps": 8, "warmup": 2, "higher_is_better": True,
                                  "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                                  "config": {"workload": "json_convnet (not in the reference; CPU restatement)"},
                                  "cpu_baseline": r, "gpu_launches": 0,
                                  "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0,
                                          "d2h_bytes_per_step": 0}}), flush=True)
            return
        bench_convnet.run(args, ClockSampler, _peaks, METRIC)
        return
    if args.impl == "reference":
        run_reference_arm(args)
        return

    import numpy as np
    import torch
    import torch.distributed as dist
    from pytorch_ddp_mnist_b200 import engine as E, ops, synth
    from pytorch_ddp_mnist_b200.dataset import ResidentDataset

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", init_method="env://", device_id=dev)
    B, K, W = args.batch, args.steps, args.warmup

    # ---- data: resident u8 dataset (same bytes on every rank), epoch-0 shard of this rank
    im, lb = synth.synth(BENCH_N, BENCH_SEED)
    resident = ResidentDataset.from_arrays(im, lb, dev)
    from torch.utils.data import DistributedSampler
    from pytorch_ddp_mnist_b200.dataset import _IndexOnly
    sampler = DistributedSampler(_IndexOnly(BENCH_N), num_replicas=world, rank=rank, shuffle=True, seed=42)
    sampler.set_epoch(0)
    shard = torch.tensor(list(iter(sampler)), dtype=torch.int64)
    need = (W + K + CHUNK + 260) * B                         # + the per-kernel timing section's 200 fresh batches
    reps = -(-need // shard.numel())
    shard_dev = shard.repeat(reps)[:need].to(dev)          # wraps after 3750/ws steps; rows long evicted by then
    shards_all_dev = None
    if world > 1:                                          # every rank's shard (a pure function of seed, epoch, rank)
        rows = []
        for r in range(world):
            sr = DistributedSampler(_IndexOnly(BENCH_N), num_replicas=world, rank=r, shuffle=True, seed=42)
            sr.set_epoch(0)
            rows.append(torch.tensor(list(iter(sr)), dtype=torch.int64).repeat(reps)[:need])
        shards_all_dev = torch.stack(rows).to(dev)
        assert torch.equal(shards_all_dev[rank], shard_dev)

    torch.manual_seed(0)
    eng = E.TrainEngine(dev, B, lr=0.01, world_size=world, rank=rank, tf32=args.tf32 and B > 512)
    eng.load_state_dict(E.init_params_like_reference())
    eng.broadcast_params()

    # ---- ws == 1: the persistent kernel runs CHUNK whole steps per launch straight off the shard.
    # ---- ws  > 1: the same, with the gradient exchange inside the kernel (fused, default); --no-fused: CUDA graph
    #      of CHUNK unrolled steps (persistent 1-step kernel -> NCCL allreduce -> flat SGD) reading a fixed index
    #      buffer that is refilled per replay.  Batches > 512 / --three-kernel: CUDA graph of the v1 kernels.
    chunk = min(CHUNK, K)
    persistent = (B <= 512) and not args.three_kernel
    idx_buf = torch.empty(chunk * B, dtype=torch.int64, device=dev)
    idx_buf.copy_(shard_dev[:chunk * B])
    batches = list(torch.split(idx_buf, B))
    for t in range(3):                                      # eager warm-up (also warms NCCL, sets func attributes)
        if persistent:
            eng.train_steps(resident, batches[t % chunk], B, 1)
        else:
            eng.train_step(resident, batches[t % chunk])
    fused = persistent and world > 1 and not args.no_fused
    if fused:
        eng.enable_fused_ddp()                              # gradient exchange inside the kernel (NVLink peer memory)
    use_graph = ((world > 1) and not fused) or not persistent
    graph = eng.capture_steps(resident, batches, persistent=persistent) if use_graph else None
    if persistent:
        kernels_per_chunk = 1 if (world == 1 or fused) else 2 * chunk   # persistent (+ sgd per step); NCCL's kernel not counted
    else:
        kernels_per_chunk = (3 if world == 1 else 4) * chunk

    def run(n_steps: int, start_step: int) -> int:
        """n_steps optimisation steps starting at global step `start_step`; returns launches."""
        done, launches = 0, 0
        while done + chunk <= n_steps:
            s = (start_step + done) * B
            if use_graph:
                idx_buf.copy_(shard_dev[s:s + chunk * B])
                graph.replay()
            else:
                ia = shards_all_dev[:, s:s + chunk * B].contiguous() if (fused and shards_all_dev is not None) else None
                eng.train_steps(resident, shard_dev[s:s + chunk * B], B, chunk, idx_all=ia)
            done += chunk
            launches += kernels_per_chunk
        if done < n_steps and not use_graph:                # remainder: one shorter launch through the very same path
            s, r = (start_step + done) * B, n_steps - done  # (so that a short warm-up also touches the peer buffers, idx_all, ...)
            ia = shards_all_dev[:, s:s + r * B].contiguous() if (fused and shards_all_dev is not None) else None
            eng.train_steps(resident, shard_dev[s:s + r * B], B, r, idx_all=ia)
            done, launches = n_steps, launches + 1
        while done < n_steps:                               # remainder of the graph path, eager
            s = (start_step + done) * B
            if persistent:
                eng.train_steps(resident, shard_dev[s:s + B], B, 1)
            else:
                eng.train_step(resident, shard_dev[s:s + B])
            done += 1
            launches += kernels_per_chunk // chunk
        return launches

    run(W, 0)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler_thr = ClockSampler(local_rank)
    sampler_thr.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    launches = run(K, W)
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    clocks = sampler_thr.stop()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = K * B * world / (ms * 1e-3)
    final_loss = eng.last_batch_loss()

    # ---- per-kernel durations, live, with CUDA events on the launching stream
    kern = _time_kernels(torch, ops, eng, resident, shard_dev, B, world, three_kernel=not persistent)

    # ---- e2e: host buffers in, loss out, every step
    e2e = _time_e2e(torch, eng, im, lb, shard, B, min(K, 2000), world, dev, dist if world > 1 else None)

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "tf32" if eng.tf32 else "f32", "data": "synthetic", "config": _config(args, {"final_batch_loss": final_loss}),
        "gpu_launches": launches, "clocks": clocks, "e2e": e2e, "roofline": kern["roofline"],
        "kernels": kern["kernels"],
    }
    if rank == 0:
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = _cpu_baseline(args)
        print(json.dumps(line), flush=True)
    if world > 1:
        # Tear-down must never hang the run: NCCL communicators captured in live CUDA graphs have
        # been seen to block in destroy_process_group().  The JSON line is already flushed, so give
        # the orderly path a few seconds and then leave.
        sys.stdout.flush()
        threading.Timer(20.0, lambda: os._exit(0)).start()
        torch.cuda.synchronize()
        dist.barrier()
        graph = None
        os._exit(0)


def _time_kernels(torch, ops, eng, resident, shard_dev, B, world, three_kernel=False):
    """Average duration of each of the step's kernels, measured with CUDA events around runs of
    that kernel alone on fresh (cold) rows, and the HBM roofline of the dominant one.
    Algorithmic bytes per launch: pixels 784 B + index 8 B (+ label 1 B) per image,
    each parameter/activation tensor the phase must read or write once."""
    hbm, which = _peaks()
    P = 473_088
    reps = 200
    lr = eng.lr if world == 1 else 0.0
    gscale = 1.0 / world
    grads = eng.grads if world > 1 else None
    params = eng.params.clone()                              # do not disturb the trained weights
    idxs = [shard_dev[i * B:(i + 1) * B] for i in range(reps)]
    mask_kw = dict(philox=(1, 0))
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]

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

    # forward = fc1_fwd + rows; backward = grads (+sgd when ws>1).  Separate them by difference.
    def fwd(idx):
        ops.step_forward(params, eng.ws, B, images=resident.images, labels=resident.labels, idx=idx,
                         prep_backward=True, loss=eng.loss, tf32=eng.tf32, **mask_kw)

    def bwd(idx):
        ops.step_backward(params, eng.ws, B, images=resident.images, idx=idx, grads=grads, grad_scale=gscale, lr=lr,
                          tf32=eng.tf32)

    def gather_only(idx):
        ops.gather_decode(resident.images, resident.labels, idx, out_x=xbuf, out_y=ybuf)

    xbuf = torch.empty(B, 784, device=params.device)
    ybuf = torch.empty(B, dtype=torch.uint8, device=params.device)
    t_fwd, t_bwd, t_gather = avg_ms(fwd), avg_ms(bwd), avg_ms(gather_only)
    bytes_fwd = B * (784 + 8 + 1) + (100_352 + 128) * 4 + 71_680 + 4 * B * 512 + B * 64   # px,idx,label + W1,b1 + W2,b2,W3 + h1d,h2,dz2,dz1 + dz3
    bytes_bwd = B * (784 + 8) + 4 * B * 512 + B * 64 + 2 * P + (P if world > 1 else 0)    # px again + activations + params r/w (+ bucket)
    bytes_gather = B * (784 + 8 + 1) + B * (3136 + 1)
    # persistent kernel: 50 steps per launch, average per step
    t_persist = None
    if B <= 512:
        nst = 50
        flat = shard_dev[:nst * B * 4]

        def persist(i):
            ops.train_steps_persistent(params, eng.ws, B, nst, images=resident.images, labels=resident.labels,
                                       idx=flat[(i % 4) * nst * B:((i % 4) + 1) * nst * B], philox=(1, 0), lr=eng.lr)
        for i in range(3):
            persist(i)
        torch.cuda.synchronize()
        ev[2].record()
        for i in range(8):
            persist(i)
        ev[3].record()
        torch.cuda.synchronize()
        t_persist = ev[2].elapsed_time(ev[3]) / (8 * nst)
    bytes_step = B * (784 + 8 + 1) + 2 * P          # pixels + index + label, parameters read once and written once
    kernels = [
        {"name": "train_persist_kernel (whole step: fwd + loss + bwd + SGD; per step of a 50-step launch)",
         "ms": t_persist, "alg_bytes": bytes_step, "gbs": (bytes_step / t_persist / 1e6) if t_persist else None},
        {"name": "step_forward (fc1_fwd_kernel + rows_kernel)", "ms": t_fwd, "alg_bytes": bytes_fwd,
         "gbs": bytes_fwd / t_fwd / 1e6},
        {"name": "step_backward (grads_kernel, SGD fused)" if world == 1 else "step_backward (grads_kernel -> bucket)",
         "ms": t_bwd, "alg_bytes": bytes_bwd, "gbs": bytes_bwd / t_bwd / 1e6},
        {"name": "gather_decode_kernel (standalone loader kernel, not in the fused step)", "ms": t_gather,
         "alg_bytes": bytes_gather, "gbs": bytes_gather / t_gather / 1e6},
    ]
    dom = kernels[0] if (t_persist and not three_kernel) else max(kernels[1:3], key=lambda k: k["ms"])
    roof = {"bound": "hbm", "achieved": dom["gbs"], "peak": hbm, "unit": "GB/s", "frac": dom["gbs"] / hbm,
            "traffic": _ncu_traffic() if dom is kernels[0] else None, "kernel": dom["name"], "peak_source": which,
            "note": "algorithmic bytes = what one step must move (DESIGN.md §4); batch-64 steps are latency-bound "
                    "(2 grid barriers + dependent round trips per step), not bandwidth-bound"}
    return {"kernels": kernels, "roofline": roof}


def _time_e2e(torch, eng, im, lb, shard, B, steps, world, dev, dist):
    """Same metric through the public step API with HOST inputs: per step, H2D of the batch's u8
    pixels + labels from pinned memory, the fused step, and a D2H read of the loss."""
    import numpy as np
    steps = max(steps, 10)
    rows = np.resize(shard.numpy(), (steps + 10) * B) % im.shape[0]      # wraps if the shard is shorter
    host_x = torch.from_numpy(im.reshape(-1, 784)[rows]).pin_memory()
    host_y = torch.from_numpy(lb[rows]).pin_memory()
    x_dev = [torch.empty(B, 784, dtype=torch.uint8, device=dev) for _ in range(2)]
    y_dev = [torch.empty(B, dtype=torch.uint8, device=dev) for _ in range(2)]
    ident = torch.arange(B, dtype=torch.int64, device=dev)
    loss_host = torch.zeros(steps + 10, 2).pin_memory()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def one(t, sync):
        b = t & 1
        x_dev[b].copy_(host_x[t * B:(t + 1) * B], non_blocking=True)
        y_dev[b].copy_(host_y[t * B:(t + 1) * B], non_blocking=True)
        eng.train_step_host_batch(x_dev[b], y_dev[b], ident)
        loss_host[t].copy_(eng.loss, non_blocking=True)
        if sync:
            torch.cuda.current_stream().synchronize()       # the reference's per-step .item() (mp:464)

    out = {}
    for name, sync in (("pipelined", False), ("sync_each_step", True)):
        for t in range(10):
            one(t, sync)
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        ev0.record()
        for t in range(10, 10 + steps):
            one(t, sync)
        ev1.record()
        torch.cuda.synchronize()
        ms = ev0.elapsed_time(ev1)
        if dist is not None:
            tt = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ms = float(tt.item())
        out[name] = steps * B * world / (ms * 1e-3)
    assert bool(torch.isfinite(loss_host[10:10 + steps, 0]).all())

    # ---- the host-fed loop (TrainEngine.train_steps_host -> ddpm_train_steps_host_fed): ONE call trains on `steps`
    # batches that sit in pinned host memory; every step's pixels + labels are copied H2D (copy stream -> device ring)
    # while the persistent kernel runs, and every step's loss is written by the kernel into pinned host memory.
    fed = None
    if B <= 512 and (world == 1 or eng.comm is not None):
        import time
        hx = host_x.view(-1, B, 784)
        hy = host_y.view(-1, B)
        hloss = torch.full((steps,), float("nan")).pin_memory()
        eng.train_steps_host(hx[:10], hy[:10], B, 10, host_loss_steps=hloss[:10])     # warm-up: ring, copy stream, func attributes
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        t0 = time.perf_counter()
        ev0.record()
        eng.train_steps_host(hx[10:10 + steps], hy[10:10 + steps], B, steps, host_loss_steps=hloss)
        ev1.record()
        torch.cuda.synchronize()
        wall_ms = (time.perf_counter() - t0) * 1e3
        ms = max(ev0.elapsed_time(ev1), wall_ms)            # the slower of the device and the host clock
        eng.check_feed()
        if dist is not None:
            tt = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ms = float(tt.item())
        assert bool(torch.isfinite(hloss).all())            # every step's loss arrived in host memory
        fed = steps * B * world / (ms * 1e-3)
    value = fed if fed is not None else out["pipelined"]
    return {"value": value, "unit": UNIT, "h2d_bytes_per_step": B * 785 + (8 if fed is not None else 0),
            "d2h_bytes_per_step": 4 if fed is not None else 8, "steps": steps,
            "api": "TrainEngine.train_steps_host (ddpm_train_steps_host_fed)" if fed is not None else
                   "TrainEngine.train_step_host_batch per step",
            "per_step_calls_value": out["pipelined"], "sync_each_step_value": out["sync_each_step"],
            "note": "value: one host-fed call for all steps - per step, B*785 B of u8 pixels + labels (+ an 8-byte sequence "
                    "word) are copied from pinned host memory into a device ring by the copy engine while the persistent "
                    "kernel runs, and the kernel stores that step's loss into pinned host memory; "
                    "per_step_calls_value: one Python call per step (2 H2D copies, 1-step launch, async D2H of the loss); "
                    "sync_each_step_value: the same with the host waiting for every loss like the reference's .item()"}


def _cpu_baseline(args):
    """The reference's CPU path (port, checked against the reference in tests/test_ref_port.py),
    2 gloo ranks on this box's cores, bounded sample."""
    code = ("import json,sys,tempfile;sys.path.insert(0,%r);"
            "from oracle import ref_port;from pytorch_ddp_mnist_b200 import synth;"
            "d=tempfile.mkdtemp();synth.write_default_files(d,60000,1000);"
            "print('CPUBASE'+json.dumps(ref_port.time_port_gloo(d,%d,%d,10,2,29890)))") % (
                REPO, args.batch, args.cpu_baseline_steps)
    try:
        out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
        r = json.loads([l for l in out.stdout.splitlines() if l.startswith("CPUBASE")][-1][7:])
        return {"value": r["images_per_s"], "unit": UNIT, "cores": r["cores"], "kind": "port",
                "sample": f"{args.cpu_baseline_steps} steps x batch {args.batch} x 2 gloo ranks ({r['threads_per_rank']} "
                          f"threads each), per-sample CDF-5 reads, loader share {r['loader_fraction']:.2f}, "
                          f"{r['seconds']:.1f} s"}
    except Exception as e:                                       # never lose the GPU line over the baseline
        return {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "port", "sample": f"failed: {e!r}"}


if __name__ == "__main__":
    main()
