"""Data-parallel parity gate: the GPU step paths against the REFERENCE's own gloo runs.

Used by `bench.py` (before its timed region, at every world size: the result travels in the JSON line
as `parity_check`) and by `tests/multigpu_check.py`.  CHECKER CODE — it compares, it is never timed.

Fixtures: `tests/golden/traj_ws{1,2,4,8}.npz`, written by `oracle/gen_golden.py` from the unmodified
`mnist_pnetcdf_cpu_mp.py` loop body (mp:439 DDP wrap, mp:453-466 loop) on 1/2/4/8 gloo ranks: per-rank sample
indices, dropout masks, losses of 20 steps (100 at ws 1), initial and final parameters.

What is run, each from the fixture's initial parameters, on this rank's fixture batches and masks:

  nccl      pre-divided flat bucket -> NCCL all-reduce(SUM) -> flat SGD between the step kernels (`--no-fused`)
  fused     the default path at ws > 1: gradient exchange inside the persistent kernel (with the dz1 row
            exchange where that is the default for this world size)
  fused_bucket   the same with the dz1 exchange switched off (plain bucket exchange)
  host_fed  the host-fed loop (`e2e` path): batches from pinned host memory, exchange inside the kernel

Bars (the same as the single-GPU trajectory tests): per-step losses within 5e-3 (max-norm,
relative to the largest loss) of the reference's over the whole run and within 1e-5 on the first three steps;
final parameters within 5e-3 (Frobenius) per tensor; every path's replicas BIT-identical across ranks; fused
paths within 1e-4 of the NCCL path.
"""
from __future__ import annotations

import os
from typing import Dict, Optional

import numpy as np
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN_WS = (1, 2, 4, 8)
TOL_TRAJ, TOL_FIRST, TOL_PATHS = 5e-3, 1e-5, 1e-4


def _replicas_identical(params: torch.Tensor, dist) -> bool:
    if dist is None:
        return True
    other = params.clone()
    dist.broadcast(other, src=0)
    same = torch.tensor([1 if torch.equal(params, other) else 0], device=params.device)
    dist.all_reduce(same, op=dist.ReduceOp.MIN)
    return bool(same.item())


def _allmax(x: float, dev, dist) -> float:
    if dist is None:
        return float(x)
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def run_gate(dev, rank: int, world: int, dist=None, steps: int = 20, include_host_fed: bool = True) -> Dict:
    """Runs every path listed above for `steps` steps and returns the `parity_check` record.  Collective: every
    rank must call it.  Never raises on a numerical miss (the record says `ok: false` and why); raises only when
    the machinery itself fails (missing library, communicator error)."""
    from pytorch_ddp_mnist_b200 import engine as E, synth
    from pytorch_ddp_mnist_b200.dataset import ResidentDataset
    keys = E.PARAM_KEYS
    have_golden = world in GOLDEN_WS
    rec: Dict = {"world_size": world, "steps": steps, "paths": {}, "ok": True, "failures": []}
    im, lb = synth.synth(synth.TRAIN_N, synth.TRAIN_SEED)
    res = ResidentDataset.from_arrays(im, lb, dev)
    pre = "" if world == 1 else f"rank{rank}."
    if have_golden:
        g = np.load(os.path.join(REPO, "tests", "golden", f"traj_ws{world}.npz"))
        steps = min(steps, g[pre + "idx"].shape[0])
        rec["steps"] = steps
        rec["golden"] = (f"tests/golden/traj_ws{world}.npz (unmodified reference loop, {world} gloo rank"
                         f"{'s' if world > 1 else ''}, oracle/gen_golden.py)")
        masks_np = np.unpackbits(g[pre + "masks"], axis=-1).astype(np.uint8)[:steps]
        idx_np = g[pre + "idx"][:steps].astype(np.int64)
        ref_losses = g[pre + "losses"][:steps].astype(np.float64)
        p0 = {k: torch.from_numpy(g[f"params0.{k}"]) for k in keys}
        full_run = steps == g[pre + "idx"].shape[0]
        idx_all_np = idx_np.reshape(1, -1) if world == 1 else \
            np.stack([g[f"rank{r}.idx"][:steps].reshape(-1) for r in range(world)]).astype(np.int64)
    else:                                           # no reference run at this world size: self-consistency only
        rec["golden"] = None
        from torch.utils.data import DistributedSampler
        from pytorch_ddp_mnist_b200.dataset import _IndexOnly
        rows = []
        for r in range(world):
            s = DistributedSampler(_IndexOnly(synth.TRAIN_N), num_replicas=world, rank=r, shuffle=True, seed=42)
            s.set_epoch(0)
            rows.append(np.array(list(iter(s)), np.int64)[:steps * 64])
        idx_all_np = np.stack(rows)
        idx_np = idx_all_np[rank].reshape(steps, 64)
        masks_np = (np.random.default_rng(500 + rank).random((steps, 64, 128)) < 0.8).astype(np.uint8)
        ref_losses, full_run = None, False
        torch.manual_seed(0)
        p0 = E.init_params_like_reference()
    masks = torch.from_numpy(masks_np).to(dev).contiguous()
    idx = torch.from_numpy(idx_np).to(dev).contiguous()
    idx_all = torch.from_numpy(np.ascontiguousarray(idx_all_np)).to(dev)
    B = idx.shape[1]

    def fail(msg):
        rec["ok"] = False
        rec["failures"].append(msg)

    def judge(name, losses, params, against: Optional[torch.Tensor]):
        """losses: np[steps] of this rank; params: flat device tensor."""
        out: Dict = {}
        if ref_losses is not None:
            e = float(np.abs(losses - ref_losses).max() / np.abs(ref_losses).max())
            e3 = float(np.abs(losses[:3] - ref_losses[:3]).max())
            out["loss_err_vs_reference"] = _allmax(e, dev, dist)
            out["first3_loss_err_vs_reference"] = _allmax(e3, dev, dist)
            if out["loss_err_vs_reference"] > TOL_TRAJ:
                fail(f"{name}: loss trajectory {out['loss_err_vs_reference']:.2e} > {TOL_TRAJ}")
            if out["first3_loss_err_vs_reference"] > TOL_FIRST:
                fail(f"{name}: first three losses {out['first3_loss_err_vs_reference']:.2e} > {TOL_FIRST}")
            if full_run:
                flat = params.detach().cpu().numpy().astype(np.float64)
                worst = 0.0
                for k, shp, off in zip(keys, E.PARAM_SHAPES, E.PARAM_OFFSETS):
                    r = g[f"paramsN.{k}"].astype(np.float64).reshape(-1)
                    worst = max(worst, float(np.linalg.norm(flat[off:off + r.size] - r) / np.linalg.norm(r)))
                out["final_params_err_vs_reference"] = worst
                if worst > TOL_TRAJ:
                    fail(f"{name}: final parameters {worst:.2e} > {TOL_TRAJ}")
        if against is not None:
            rel = float((params - against).abs().max() / against.abs().max())
            out["params_err_vs_nccl_path"] = _allmax(rel, dev, dist)
            if out["params_err_vs_nccl_path"] > TOL_PATHS:
                fail(f"{name}: differs from the NCCL path by {out['params_err_vs_nccl_path']:.2e} > {TOL_PATHS}")
        out["replicas_identical"] = _replicas_identical(params, dist)
        if not out["replicas_identical"]:
            fail(f"{name}: replicas diverged")
        out["finite"] = bool(np.isfinite(losses).all())
        if not out["finite"]:
            fail(f"{name}: non-finite loss")
        rec["paths"][name] = out

    def fresh():
        e = E.TrainEngine(dev, B, world_size=world, rank=rank)
        if world > 1 and rank != 0:                 # ranks start different; the DDP wrap makes rank 0's win (mp:439)
            torch.manual_seed(100 + rank)
            e.load_state_dict(E.init_params_like_reference())
        else:
            e.load_state_dict(p0)
        e.broadcast_params(0)
        return e

    # ---- NCCL between kernels (ws == 1: the plain persistent kernel, one step per launch)
    eng = fresh()
    ls = []
    for t in range(steps):
        eng.train_steps(res, idx[t], B, 1, masks=masks[t:t + 1])
        ls.append(eng.last_batch_loss())
    base_name = "nccl" if world > 1 else "persistent_1step"
    judge(base_name, np.array(ls, np.float64), eng.params, None)
    base = eng.params.clone()

    def resident_run(name, env_off: bool):
        e = fresh()
        if world > 1:
            e.enable_fused_ddp()
        out = torch.zeros(steps, device=dev)
        old = os.environ.get("DDPM_NO_ACT_EXCHANGE")
        if env_off:
            os.environ["DDPM_NO_ACT_EXCHANGE"] = "1"
        try:
            e.train_steps(res, idx.reshape(-1), B, steps, masks=masks, loss_steps=out,
                          idx_all=idx_all if world > 1 else None)
            torch.cuda.synchronize(dev)
        finally:
            if env_off:
                if old is None:
                    os.environ.pop("DDPM_NO_ACT_EXCHANGE", None)
                else:
                    os.environ["DDPM_NO_ACT_EXCHANGE"] = old
        e.check_comm()
        judge(name, out.cpu().numpy().astype(np.float64), e.params, base)
        return e

    if world > 1:
        resident_run("fused", env_off=False)        # the default exchange of this world size
        twin = resident_run("fused_bucket", env_off=True)
    else:
        twin = resident_run("persistent_multistep", env_off=False)

    if include_host_fed:
        e = fresh()
        if world > 1:
            e.enable_fused_ddp()
        flat = idx_np.reshape(-1)
        hx = torch.from_numpy(im.reshape(-1, 784)[flat]).reshape(steps, B, 784).contiguous().pin_memory()
        hy = torch.from_numpy(lb[flat]).reshape(steps, B).contiguous().pin_memory()
        hl = torch.full((steps,), float("nan")).pin_memory()
        e.train_steps_host(hx, hy, B, steps, masks=masks, host_loss_steps=hl, slots=16)
        torch.cuda.synchronize(dev)
        e.check_comm()
        e.check_feed()
        judge("host_fed", hl.numpy().astype(np.float64), e.params, base)
        # same arithmetic as the resident loop with the plain bucket exchange: bit for bit
        same = torch.tensor([1 if torch.equal(e.params, twin.params) else 0], device=dev)
        if dist is not None:
            dist.all_reduce(same, op=dist.ReduceOp.MIN)
        rec["paths"]["host_fed"]["bit_identical_to_resident_loop"] = bool(same.item())
        if not bool(same.item()):
            fail("host_fed: differs from the resident loop on the same batches")

    errs = [p.get("loss_err_vs_reference") for p in rec["paths"].values() if "loss_err_vs_reference" in p]
    rec["max_loss_err"] = max(errs) if errs else None
    rec["replicas_identical"] = all(p["replicas_identical"] for p in rec["paths"].values())
    rec["tolerances"] = {"loss_trajectory_rel": TOL_TRAJ, "first3_abs": TOL_FIRST, "paths_rel": TOL_PATHS}
    if dist is not None:                            # every rank reports the same verdict
        ok = torch.tensor([1 if rec["ok"] else 0], device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        rec["ok"] = bool(ok.item())
    del res
    return rec
