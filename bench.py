#!/usr/bin/env python
"""Benchmark of the hot path: one data-parallel MNIST training step (reference model `ref_mlp`,
`mnist_pnetcdf_cpu_mp.py:453-466`) on synthetic MNIST-shaped data.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch 64] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

Prints ONE JSON line on rank 0.  DESIGN.md §4 says what every field means.

How the headline number is timed (round 2): the K-step region is run once the literal way (barrier +
synchronize, CUDA events around K steps: `single_shot`) and then `repeats` more times back to back, each
repetition bracketed by its own CUDA events and - at N > 1 - started from an on-device rendezvous (a 4-byte NCCL
all-reduce enqueued right before the start event, so that no rank's K-step launch is charged with another rank's
late exit from Python).  `value` = K*B*N / median over repetitions of (max over ranks of the repetition's time);
min / median / max and the single-shot figure are in `timing`.  At the driver's `--steps 20` one repetition is ONE
230-microsecond kernel launch: a single sample of it is noise.
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
CHUNK = 100                # steps per launch of the persistent kernel (or unrolled per CUDA graph)
NCU_SUMMARY = os.path.join("profiles", "r02_persist_ncu_full_summary.csv")


def _ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per STEP of train_persist_kernel, from the committed
    `ncu --set full` capture of the kernel this build ships (profiles/README.md says which commit and launch);
    (None, why) if the capture is absent."""
    import csv
    p = os.path.join(REPO, NCU_SUMMARY)
    try:
        rows = list(csv.reader(open(p)))
        hdr, units = rows[0], rows[1]
        mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        # the capture holds launches of different step counts: traffic per step = slope between them
        pts = []
        for r in rows[2:]:
            tot = 0.0
            for name in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                i = hdr.index(name)
                tot += float(r[i]) * mult[units[i]]
            steps = int(r[hdr.index("launch__steps")]) if "launch__steps" in hdr else 1
            pts.append((steps, tot))
        if len(pts) >= 2 and pts[-1][0] != pts[0][0]:
            per_step = (pts[-1][1] - pts[0][1]) / (pts[-1][0] - pts[0][0])
        else:
            per_step = pts[0][1] / pts[0][0]
        return per_step, NCU_SUMMARY
    except Exception as e:
        return None, f"no capture of this build ({type(e).__name__})"


def _peaks():
    p = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ------------------------------------------------------------------------------------ clocks
class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons of one GPU through NVML while the timed region runs; `mark()` takes one
    sample synchronously (called right before and right after the region, so short regions are covered too)."""

    def __init__(self, index: int, period_s: float = 0.002):
        super().__init__(daemon=True)
        self.index, self.period = index, period_s
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop_evt = threading.Event()
        self._lock = threading.Lock()
        self.err = None
        self._h = self._nv = None
        try:
            import pynvml as nv
            nv.nvmlInit()
            self._nv = nv
            self._h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(self._h, nv.NVML_CLOCK_SM)
            self._names = {
                nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
                nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: "hw_power_brake",
            }
        except Exception as e:                                   # NVML missing: report it
            self.err = repr(e)

    def mark(self):
        if self._h is None:
            return
        try:
            nv = self._nv
            mhz = nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)
            r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h)
            with self._lock:
                self.samples.append(mhz)
                for bit, name in self._names.items():
                    if r & bit:
                        self.reasons.add(name)
        except Exception as e:
            self.err = repr(e)

    def run(self):
        while not self._stop_evt.is_set() and self._h is not None:
            self.mark()
            time.sleep(self.period)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s), **({"error": self.err} if self.err else {})}


# ------------------------------------------------------------------------------------ reference arm
def run_reference_arm(args) -> None:
    """`--impl reference`: the reference's own CPU path (gloo DDP, per-sample netCDF loader) on the host cores, via
    oracle/ref_port.py (the Python reference cannot travel to the GPU box; the port is checked against it in
    tests/test_ref_port.py).  World size: the reference's CPU run is configs[0] = 2 gloo ranks; under `--gpus N`
    it runs max(2, N) gloo ranks on the same cores (`mnist_pnetcdf_cpu_mp.py:213-254, 439`), so that every point of
    the scaling run has the same-N reference beside it.  Under torchrun only rank 0 works; the port's workers drop
    the launcher's environment and rendezvous through a private FileStore (they used to inherit
    TORCHELASTIC_USE_AGENT_STORE and wait for a store nobody served)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import ref_port
    from pytorch_ddp_mnist_b200 import synth
    ref_ws = max(2, args.gpus)
    steps = max(1, min(args.steps, 400))            # bounded sample: ~10 ms/step on CPU
    warmup = max(1, min(args.warmup, 20))
    with tempfile.TemporaryDirectory() as d:
        synth.write_default_files(d, 60_000, 1_000)
        r = ref_port.time_port_gloo(d, args.batch, steps, warmup, world_size=ref_ws, timeout_s=420.0)
    v = r["images_per_s"]
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": 1e3 * r["seconds"] / steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "model": "ref_mlp",
        "config": _config(args, extra={"reference_world_size": ref_ws}),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": r["cores"], "kind": "port",
                         "sample": f"{steps} steps x batch {args.batch} x {ref_ws} gloo ranks "
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
         "step_kernel": "three-kernel (v1)" if args.three_kernel or args.batch > 512 else "persistent",
         "gradient_exchange": "NCCL allreduce between kernels" if args.no_fused else
                              "in-kernel over NVLink peer memory (flag-in-data pushes, sums in fixed rank order), ws > 1 only"}
    if args.batch > 512:
        c["tensor_cores"] = "fc1 forward + dW1 on tcgen05 TF32" if args.tf32 else "off (fp32 FFMA)"
    if extra:
        c.update(extra)
    return c


def _median(xs):
    s = sorted(xs)
    n = len(s)
    return s[n // 2] if n % 2 else 0.5 * (s[n // 2 - 1] + s[n // 2])


# ------------------------------------------------------------------------------------ our arm
def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=200)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--repeats", type=int, default=0, help="repetitions of the K-step region (0 = auto: 30 for K <= 200, "
                                                            "fewer for long regions, never below 5)")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity-gate", action="store_true", help="skip the pre-run parity gate (tests/parity_gate.py)")
    ap.add_argument("--no-context", action="store_true", help="skip the context numbers (launch floor, NCCL latency, torch eager, ConvNet sub-record)")
    ap.add_argument("--no-fused", action="store_true", help="ws > 1: NCCL allreduce between kernels instead of the fused peer-memory exchange")
    ap.add_argument("--three-kernel", action="store_true", help="use the v1 three-kernel step instead of the persistent kernel")
    ap.add_argument("--tf32", action="store_true", help="batch > 512 only: fc1 forward and dW1 on tcgen05 TF32 tensor cores (fp32 accumulate)")
    ap.add_argument("--cpu-baseline-steps", type=int, default=150)
    ap.add_argument("--model", default="ref_mlp", choices=["ref_mlp", "json_convnet"],
                    help="ref_mlp = the model the reference defines (default, the graded path); json_convnet = the ConvNet "
                         "BASELINE.json names, which is not in the reference (bench_convnet.py)")
    ap.add_argument("--fp32", action="store_true", help="json_convnet only: fp32 FFMA kernels instead of tcgen05 TF32")
    ap.add_argument("--config", default="default", choices=["default", "scaled6m"],
                    help="scaled6m = BASELINE.json configs[4]: 6 M-sample CDF-5 file, loader + one full epoch (bench_scaled.py)")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.config == "scaled6m":
        import bench_scaled
        bench_scaled.run(args, ClockSampler, METRIC)
        return
    if args.model == "json_convnet":
        import bench_convnet
        if args.impl == "reference":            # no reference code exists for this model: the CPU arm is the restatement
            if int(os.environ.get("RANK", "0")) == 0:
                r = bench_convnet._cpu_baseline(args.batch, 150)
                print(json.dumps({"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT,
                                  "n_gpus": args.gpus, "steps": 8, "warmup": 2, "higher_is_better": True,
                                  "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                                  "model": "json_convnet",
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
    R = args.repeats if args.repeats > 0 else max(5, min(30, 6000 // max(K, 1)))

    # ---- parity gate: every step path of this world size against the reference's own gloo run, BEFORE timing
    parity = None
    if not args.no_parity_gate and B <= 512:
        sys.path.insert(0, os.path.join(REPO, "tests"))
        import parity_gate
        try:
            parity = parity_gate.run_gate(dev, rank, world, dist if world > 1 else None)
        except Exception as e:                               # the gate's machinery failed: say so, keep the line
            parity = {"ok": False, "failures": [f"gate raised {e!r}"], "world_size": world}
        torch.cuda.synchronize()

    # ---- data: resident u8 dataset (same bytes on every rank), epoch-0 shard of this rank
    im, lb = synth.synth(BENCH_N, BENCH_SEED)
    resident = ResidentDataset.from_arrays(im, lb, dev)
    from torch.utils.data import DistributedSampler
    from pytorch_ddp_mnist_b200.dataset import _IndexOnly
    sampler = DistributedSampler(_IndexOnly(BENCH_N), num_replicas=world, rank=rank, shuffle=True, seed=42)
    sampler.set_epoch(0)
    shard = torch.tensor(list(iter(sampler)), dtype=torch.int64)
    need = (W + (R + 1) * K + CHUNK + 260) * B               # + the per-kernel timing section's 200 fresh batches
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

    def allmax(vals):
        if world == 1:
            return [float(v) for v in vals]
        t = torch.tensor(list(vals), dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(v) for v in t.tolist()]

    run(W, 0)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler_thr = ClockSampler(local_rank)
    sampler_thr.start()
    sampler_thr.mark()
    # ---- (1) the literal region: barrier + synchronize, K steps, synchronize
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    launches_per_rep = run(K, W)
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ms_single = allmax([e0.elapsed_time(e1)])[0]
    # ---- (2) R repetitions of the same K-step region, back to back, each with its own events; at ws > 1 each one
    #      starts from an on-device rendezvous (tiny NCCL all-reduce on the same stream)
    tick = torch.zeros(1, device=dev)
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(R)]
    for r in range(R):
        if world > 1:
            dist.all_reduce(tick)
        evs[r][0].record()
        run(K, W + (r + 1) * K)
        evs[r][1].record()
    torch.cuda.synchronize()
    sampler_thr.mark()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    clocks = sampler_thr.stop()
    rep_ms = allmax([a.elapsed_time(b) for a, b in evs])       # per repetition: max over ranks
    ms = _median(rep_ms)
    value = K * B * world / (ms * 1e-3)
    launches = launches_per_rep * (R + 1)
    final_loss = eng.last_batch_loss()
    timing = {"repeats": R, "steps_per_repeat": K, "us_per_step_min": 1e3 * min(rep_ms) / K,
              "us_per_step_median": 1e3 * ms / K, "us_per_step_max": 1e3 * max(rep_ms) / K,
              "single_shot_us_per_step": 1e3 * ms_single / K,
              "single_shot_value": K * B * world / (ms_single * 1e-3),
              "rendezvous": "4-byte NCCL all-reduce on the stream before each repetition's start event" if world > 1 else None,
              "note": "value = median repetition; single_shot = one K-step region after barrier + synchronize, which at "
                      "N > 1 also contains the ranks' skew leaving the barrier (real for an epoch that is one launch)"}

    # ---- roofline of the step kernel FROM THE TIMED REGION (a launch = `chunk` steps; per step)
    hbm, which = _peaks()
    P = 473_088
    bytes_step = B * (784 + 8 + 1) + 2 * P          # pixels + index + label, parameters read once and written once
    step_s = ms * 1e-3 / K
    traffic, traffic_src = _ncu_traffic() if persistent else (None, "n/a")
    roof = {"bound": "hbm", "achieved": bytes_step / step_s / 1e9, "peak": hbm, "unit": "GB/s",
            "frac": bytes_step / step_s / 1e9 / hbm, "traffic": traffic, "traffic_source": traffic_src,
            "kernel": "train_persist_kernel (whole step: fwd + loss + bwd + SGD), per step of the timed launches"
                      if persistent else "three-kernel step (fc1_fwd + rows + grads), per step of the timed region",
            "alg_bytes_per_step": bytes_step, "peak_source": which,
            "note": "algorithmic bytes = what one step must move (DESIGN.md §3.1); batch-64 steps are latency-bound "
                    "(dependent L2 round trips per step), not bandwidth-bound"}

    # ---- per-kernel durations of the three-kernel path + standalone loader kernel, live, CUDA events
    kern = _time_kernels(torch, ops, eng, resident, shard_dev, B, world)

    # ---- e2e: host buffers in, loss out, every step
    e2e = _time_e2e(torch, eng, im, lb, shard, B, 2000, world, dev, dist if world > 1 else None)

    context = None
    if not args.no_context:
        context = _context_numbers(torch, ops, eng, dev, B, world, dist if world > 1 else None, rank)

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "tf32" if eng.tf32 else "f32", "data": "synthetic", "model": "ref_mlp",
        "config": _config(args, {"final_batch_loss": final_loss}),
        "gpu_launches": launches, "clocks": clocks, "timing": timing, "e2e": e2e, "roofline": roof,
        "kernels": kern, "parity_check": parity, "context": context,
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


def _time_kernels(torch, ops, eng, resident, shard_dev, B, world):
    """Average duration of the three-kernel path's kernels and of the standalone loader kernel, measured with CUDA
    events around runs of that kernel alone on fresh (cold) rows.  Algorithmic bytes per launch:
    pixels 784 B + index 8 B (+ label 1 B) per image, each parameter/activation tensor the phase reads or writes."""
    P = 473_088
    reps = 200
    lr = eng.lr if world == 1 else 0.0
    gscale = 1.0 / world
    grads = eng.grads if world > 1 else None
    params = eng.params.clone()                              # do not disturb the trained weights
    idxs = [shard_dev[i * B:(i + 1) * B] for i in range(reps)]
    mask_kw = dict(philox=(1, 0))
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
    return [
        {"name": "step_forward (fc1_fwd_kernel + rows_kernel; the autograd / eval / batch > 512 route)", "ms": t_fwd,
         "alg_bytes": bytes_fwd, "gbs": bytes_fwd / t_fwd / 1e6},
        {"name": "step_backward (grads_kernel, SGD fused)" if world == 1 else "step_backward (grads_kernel -> bucket)",
         "ms": t_bwd, "alg_bytes": bytes_bwd, "gbs": bytes_bwd / t_bwd / 1e6},
        {"name": "gather_decode_kernel (standalone loader kernel, not in the fused step)", "ms": t_gather,
         "alg_bytes": bytes_gather, "gbs": bytes_gather / t_gather / 1e6},
    ]


def _context_numbers(torch, ops, eng, dev, B, world, dist, rank):
    """Numbers that put the step time in context (SURVEY.md §8(d) "Timing"), none of them part of `value`:
    the empty cooperative-launch floor of the step kernel's launch shape, the standalone latency of the gradient
    exchange (NCCL all-reduce of the 473,088-byte bucket, and the library's own in-kernel exchange), a stock
    torch-eager step of the same model on the same GPU, and BASELINE.json's ConvNet through this library."""
    out = {}
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]

    def timed(fn, n, warm=5):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        ev[0].record()
        for _ in range(n):
            fn()
        ev[1].record()
        torch.cuda.synchronize()
        ms = ev[0].elapsed_time(ev[1]) / n
        if dist is not None:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    try:
        out["empty_cooperative_launch_us"] = 1e3 * timed(lambda: ops.launch_floor(1), 200)
        out["empty_cooperative_launch_note"] = ("an empty kernel with the step kernel's grid (148 x 256 threads, same "
                                                "dynamic shared memory), launched cooperatively, back to back")
    except Exception as e:                                    # context only: never lose the line over it
        out["empty_cooperative_launch_us"] = None
        out["empty_cooperative_launch_note"] = repr(e)
    if dist is not None:
        bucket = torch.zeros(118_272, device=dev)
        ms = timed(lambda: dist.all_reduce(bucket), 200)
        out["nccl_allreduce_bucket_us"] = 1e3 * ms
        out["nccl_allreduce_bus_gbs"] = 473_088 / (ms * 1e-3) * 2 * (world - 1) / world / 1e9
        if eng.comm is not None:
            try:
                n = 200
                ms = timed(lambda: ops.exchange_only(eng.comm, n), 3, warm=1) / n
                out["fused_exchange_only_us"] = 1e3 * ms
                out["fused_exchange_bus_gbs"] = 473_088 / (ms * 1e-3) * 2 * (world - 1) / world / 1e9
                out["fused_exchange_note"] = ("the step kernel's gradient exchange alone: every CTA exchanges its share of "
                                              "the 473,088-byte bucket per step, no forward/backward work; bus GB/s = "
                                              "bytes/time x 2(n-1)/n against 900 GB/s per direction")
            except Exception as e:
                out["fused_exchange_only_us"] = None
                out["fused_exchange_note"] = repr(e)
    # stock torch eager on the same GPU (sanity ceiling, not graded): same model, same batch, fp32, no TF32
    try:
        from torch import nn
        prev = torch.backends.cuda.matmul.allow_tf32
        torch.backends.cuda.matmul.allow_tf32 = False
        m = nn.Sequential(nn.Linear(784, 128), nn.ReLU(), nn.Dropout(0.2), nn.Linear(128, 128), nn.ReLU(),
                          nn.Linear(128, 10, bias=False)).to(dev)
        opt = torch.optim.SGD(m.parameters(), lr=0.01)
        lossf = nn.CrossEntropyLoss()
        px = torch.randint(0, 256, (4096, 784), dtype=torch.uint8, device=dev)
        lab = torch.randint(0, 10, (4096,), dtype=torch.int64, device=dev)
        ii = torch.randint(0, 4096, (B,), device=dev)

        def eager():
            x = ((px[ii].float() / 255.0) - 0.1307) / 0.3081
            opt.zero_grad(set_to_none=True)
            l = lossf(m(x), lab[ii])
            l.backward()
            opt.step()
        # single process: the other ranks idle (no DDP wrap) - this is a per-GPU figure
        for _ in range(10):
            eager()
        torch.cuda.synchronize()
        ev[0].record()
        for _ in range(200):
            eager()
        ev[1].record()
        torch.cuda.synchronize()
        ms = ev[0].elapsed_time(ev[1]) / 200
        out["torch_eager_step_us"] = 1e3 * ms
        out["torch_eager_images_per_s_per_gpu"] = B / (ms * 1e-3)
        out["torch_eager_note"] = "stock torch 2.11 eager (cuBLAS/ATen kernels), resident u8 pixels, same model/batch, one GPU, no DDP"
        torch.backends.cuda.matmul.allow_tf32 = prev
    except Exception as e:
        out["torch_eager_step_us"] = None
        out["torch_eager_note"] = repr(e)
    # BASELINE.json's ConvNet (not in the reference): a short resident-data measurement so that the driver's record
    # carries it; `bench.py --model json_convnet` is the full line
    try:
        import bench_convnet
        out["json_convnet"] = bench_convnet.quick_record(torch, dev, 64, world, rank, dist)
    except Exception as e:
        out["json_convnet"] = {"error": repr(e)}
    return out


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
    psteps = min(steps, 300)
    for name, sync in (("pipelined", False), ("sync_each_step", True)):
        for t in range(10):
            one(t, sync)
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        ev0.record()
        for t in range(10, 10 + psteps):
            one(t, sync)
        ev1.record()
        torch.cuda.synchronize()
        ms = ev0.elapsed_time(ev1)
        if dist is not None:
            tt = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ms = float(tt.item())
        out[name] = psteps * B * world / (ms * 1e-3)
    assert bool(torch.isfinite(loss_host[10:10 + psteps, 0]).all())

    # ---- the host-fed loop (TrainEngine.train_steps_host -> ddpm_train_steps_host_fed): ONE call trains on `steps`
    # batches that sit in pinned host memory; every step's pixels + labels are copied H2D (copy stream -> device ring)
    # while the persistent kernel runs, and every step's loss is written by the kernel into pinned host memory.
    fed = None
    if B <= 512 and (world == 1 or eng.comm is not None):
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
            "print('CPUBASE'+json.dumps(ref_port.time_port_gloo(d,%d,%d,10,2,timeout_s=300.0)))") % (
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
