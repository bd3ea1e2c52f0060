"""The training-step engine: flat parameter bucket + fused CUDA step + (for ws > 1) one allreduce.

Replaces the loop body `mp:455-465` of `mnist_pnetcdf_cpu_mp.py`:

    x.view -> optimizer.zero_grad() -> model(x) -> loss(y_hat, y) -> backward() [DDP allreduce]
    -> optimizer.step() -> epoch_loss += batch_loss.item() / B

with, per step, on one stream:

    ws == 1:  G1(fc1, decode fused) -> ROWS(fwd+loss+bwd chain) -> GRADS(+SGD fused)
    ws  > 1:  G1 -> ROWS -> GRADS(bucket <- g/ws) -> allreduce(bucket, SUM) -> SGD(flat)

Gradients live in ONE contiguous 473,088-byte bucket (the reference's DDP has exactly one bucket
of that size), pre-divided by world_size as the c10d reducer does, so the collective is a plain SUM.
The loss stays on the device: `loss[0]` is the batch mean, `loss[1]` accumulates the reference's
`epoch_loss += batch_loss / B` bookkeeping; the host reads it once per epoch.
"""
from __future__ import annotations

from typing import Callable, Dict, Iterable, List, Optional, Sequence, Tuple

import torch

from . import ops
from ._lib import N_PARAMS, OFF_B1, OFF_B2, OFF_W1, OFF_W2, OFF_W3

PARAM_KEYS = ("0.weight", "0.bias", "3.weight", "3.bias", "5.weight")    # state_dict order, mp:412-422
PARAM_SHAPES = ((128, 784), (128,), (128, 128), (128,), (10, 128))
PARAM_OFFSETS = (OFF_W1, OFF_B1, OFF_W2, OFF_B2, OFF_W3)


def init_params_like_reference(generator: Optional[torch.Generator] = None) -> Dict[str, torch.Tensor]:
    """nn.Linear's default init (kaiming_uniform(a=sqrt(5)) == U(-1/sqrt(in), 1/sqrt(in)) for both
    weight and bias), drawn in `create_model()`'s order (mp:412-422) so that, with the same
    torch.manual_seed, the values equal the reference's."""
    from torch import nn
    if generator is not None:
        raise NotImplementedError("nn.Linear draws from the global generator; seed it with torch.manual_seed")
    m = nn.Sequential(nn.Linear(28 * 28, 128), nn.ReLU(), nn.Dropout(0.2), nn.Linear(128, 128), nn.ReLU(),
                      nn.Linear(128, 10, bias=False))
    return {k: v.detach().clone() for k, v in m.state_dict().items()}


class TrainEngine:
    def __init__(self, device, max_batch: int, lr: float = 0.01, world_size: int = 1, rank: int = 0,
                 group=None, seed: int = 42, tf32: bool = False):
        self.device = torch.device(device)
        self.tf32 = bool(tf32)                    # fc1 forward / dW1 on tcgen05 TF32 (three-kernel step, batch >= 128 / 2048)
        self.max_batch = int(max_batch)
        self.lr = float(lr)
        self.world_size, self.rank, self.group = int(world_size), int(rank), group
        self.seed = int(seed) + 0x9E3779B97F4A7C15 * self.rank & 0xFFFFFFFFFFFFFFFF
        self.params = torch.zeros(N_PARAMS, dtype=torch.float32, device=self.device)
        self.grads = torch.zeros(N_PARAMS, dtype=torch.float32, device=self.device)   # the bucket
        self.loss = torch.zeros(2, dtype=torch.float32, device=self.device)
        self.ws = ops.Workspace(self.max_batch, self.device)
        self.step_count = 0                       # Philox stream position (perf-mode dropout)
        self.comm = None                          # fused data-parallel communicator (enable_fused_ddp)
        import os as _os
        # two-piece bucket at batch >= 1024, ws > 1 (the small piece's all-reduce overlaps fc1's weight-gradient kernel).
        # Off by default: measured at batch 4096 it LOSES - 2 GPUs 121.6 vs 113.5 us/step, 8 GPUs 146.6 vs 129.4 - two
        # NCCL launches and one more kernel cost more than the 71 KB all-reduce they hide (profiles/r02_b4096_*.json)
        self.overlap_allreduce = _os.environ.get("DDPM_OVERLAP_ALLREDUCE") is not None
        self._feeds = {}                          # (batch, slots) -> ops.HostFeed (train_steps_host)

    # ------------------------------------------------------------------ parameters
    def views(self, flat: Optional[torch.Tensor] = None) -> Dict[str, torch.Tensor]:
        flat = self.params if flat is None else flat
        out = {}
        for k, shp, off in zip(PARAM_KEYS, PARAM_SHAPES, PARAM_OFFSETS):
            n = 1
            for s in shp:
                n *= s
            out[k] = flat[off:off + n].view(shp)
        return out

    def load_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        v = self.views()
        for k in PARAM_KEYS:
            v[k].copy_(sd[k].to(torch.float32))

    def state_dict(self) -> Dict[str, torch.Tensor]:
        """CPU tensors under the reference's keys (what `torch.save(model.state_dict())` holds, mp:511)."""
        return {k: t.detach().cpu().clone() for k, t in self.views().items()}

    def broadcast_params(self, src: int = 0) -> None:
        """DDP construction semantics (mp:439): every rank adopts rank `src`'s parameters."""
        if self.world_size > 1:
            import torch.distributed as dist
            dist.broadcast(self.params, src=src, group=self.group)

    def enable_fused_ddp(self, strict: bool = True) -> bool:
        """Switch the ws > 1 path from `kernel -> NCCL allreduce -> SGD kernel` per step to the fused
        kernel (gradient exchange over NVLink peer memory inside the persistent step kernel).  Collective.
        The fused exchange needs all ranks on one host, world <= 8 and peer access between every pair of GPUs;
        with `strict=False` a job that cannot have it (multi-node SLURM/MPI wire-ups, > 8 ranks, no P2P) keeps the
        NCCL path on every rank and this returns False after a one-line notice on rank 0."""
        if self.world_size <= 1 or self.comm is not None:
            return self.comm is not None
        ok, why = ops.FusedComm.usable(self.world_size, self.group)
        if ok:
            try:
                self.comm = ops.FusedComm(self.rank, self.world_size, self.group)
                return True
            except ops._lib.DDPMError as e:               # agreed on by all ranks inside FusedComm
                ok, why = False, str(e)
        if strict:
            raise RuntimeError(f"fused data-parallel step unavailable: {why}")
        if self.rank == 0:
            print(f"pytorch_ddp_mnist_b200: in-kernel gradient exchange unavailable ({why}); "
                  f"using NCCL all-reduce between kernels")
        return False

    def check_comm(self) -> None:
        """Blocking.  Raises if a wait inside the step kernel gave up (a peer that never showed up, ~3 s bound)."""
        if self.comm is not None and self.comm.error():
            raise RuntimeError("fused data-parallel step: a peer did not show up within the exchange timeout")
        if self.ws.error():
            raise RuntimeError("persistent step kernel: a wait for another block's data timed out (results invalid)")

    # ------------------------------------------------------------------ steps
    def train_step(self, resident, idx: torch.Tensor, mask: Optional[torch.Tensor] = None,
                   philox_step: Optional[int] = None) -> None:
        """Enqueue one optimisation step on the current stream for the samples `idx` (i64, device).
        Dropout: `mask` u8[B,128] (parity mode) or the Philox stream at `philox_step`
        (default: an internal counter)."""
        b = idx.numel()
        if mask is None:
            st = self.step_count if philox_step is None else philox_step
            kw = dict(philox=(self.seed, st))
        else:
            kw = dict(mask=mask)
        self.step_count += 1
        ops.step_forward(self.params, self.ws, b, images=resident.images, labels=resident.labels, idx=idx,
                         prep_backward=True, loss=self.loss, tf32=self.tf32, **kw)
        if self.world_size == 1:
            ops.step_backward(self.params, self.ws, b, images=resident.images, idx=idx, lr=self.lr, tf32=self.tf32)
        else:
            import torch.distributed as dist
            if b >= 1024 and self.overlap_allreduce:
                # the bucket leaves in two pieces, as DDP's hooks let it leave during backward (mp:439, 462): everything
                # behind fc1's weights (71,680 B) is all-reduced while fc1's weight gradient is still being computed
                small, big = self.grads[OFF_B1:], self.grads[:OFF_B1]
                ops.step_backward(self.params, self.ws, b, images=resident.images, idx=idx, grads=self.grads,
                                  grad_scale=1.0 / self.world_size, tf32=self.tf32, part="small")
                w_small = dist.all_reduce(small, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
                ops.step_backward(self.params, self.ws, b, images=resident.images, idx=idx, grads=self.grads,
                                  grad_scale=1.0 / self.world_size, tf32=self.tf32, part="w1")
                w_big = dist.all_reduce(big, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
                w_small.wait()
                w_big.wait()
            else:
                ops.step_backward(self.params, self.ws, b, images=resident.images, idx=idx, grads=self.grads,
                                  grad_scale=1.0 / self.world_size, tf32=self.tf32)
                dist.all_reduce(self.grads, op=dist.ReduceOp.SUM, group=self.group)
            ops.sgd_step(self.params, self.grads, self.lr)

    def train_steps(self, resident, idx_flat: torch.Tensor, batch: int, nsteps: int,
                    masks: Optional[torch.Tensor] = None, philox_step0: Optional[int] = None,
                    loss_steps: Optional[torch.Tensor] = None, idx_all: Optional[torch.Tensor] = None) -> None:
        """`nsteps` consecutive optimisation steps over `idx_flat` (i64[nsteps*batch], device).
        batch <= 512: the persistent kernel — ONE launch for all steps at ws == 1; at ws > 1 one
        launch per step writes the pre-divided bucket, then allreduce + flat SGD.  Larger batches
        fall back to the three-kernel step.  `idx_all` i64[world, nsteps*batch] (every rank's batches, row
        `rank` == idx_flat) lets the fused kernel replace fc1's gradient exchange by the dz1 exchange."""
        from ._lib import PERSIST_MAX_BATCH
        st0 = self.step_count if philox_step0 is None else philox_step0
        self.step_count += nsteps
        if batch > PERSIST_MAX_BATCH:
            for t in range(nsteps):
                self.step_count -= 1
                self.train_step(resident, idx_flat[t * batch:(t + 1) * batch],
                                mask=None if masks is None else masks[t], philox_step=st0 + t)
            return
        kw = dict(images=resident.images, labels=resident.labels)
        if self.world_size == 1:
            mk = dict(masks=masks) if masks is not None else dict(philox=(self.seed, st0))
            ops.train_steps_persistent(self.params, self.ws, batch, nsteps, idx=idx_flat, lr=self.lr,
                                       loss=self.loss, loss_steps=loss_steps, **mk, **kw)
            return
        if self.comm is not None:
            mk = dict(masks=masks) if masks is not None else dict(philox=(self.seed, st0))
            ops.train_steps_fused_ddp(self.comm, self.params, self.ws, batch, nsteps, idx=idx_flat, idx_all=idx_all,
                                      lr=self.lr, loss=self.loss, loss_steps=loss_steps, **mk, **kw)
            return
        import torch.distributed as dist
        for t in range(nsteps):
            mk = dict(masks=masks[t:t + 1]) if masks is not None else dict(philox=(self.seed, st0 + t))
            ops.train_steps_persistent(self.params, self.ws, batch, 1, idx=idx_flat[t * batch:(t + 1) * batch],
                                       grads=self.grads, grad_scale=1.0 / self.world_size, lr=0.0, loss=self.loss,
                                       loss_steps=None if loss_steps is None else loss_steps[t:t + 1], **mk, **kw)
            dist.all_reduce(self.grads, op=dist.ReduceOp.SUM, group=self.group)
            ops.sgd_step(self.params, self.grads, self.lr)

    def train_steps_host(self, host_pixels: torch.Tensor, host_labels: torch.Tensor, batch: int, nsteps: int,
                         masks: Optional[torch.Tensor] = None, philox_step0: Optional[int] = None,
                         host_loss_steps: Optional[torch.Tensor] = None, slots: Optional[int] = None) -> None:
        """The reference's inner loop with a host-side loader (mp:452-466) as one call: `nsteps` batches in
        PINNED host memory (u8 [nsteps, batch, 784] + u8 [nsteps, batch]) are streamed into a device ring
        while the persistent kernel trains on them; every step's loss lands in `host_loss_steps` (pinned
        f32[nsteps]).  ws > 1: needs `enable_fused_ddp()` (gradient average inside the kernel)."""
        from ._lib import PERSIST_MAX_BATCH
        if batch > PERSIST_MAX_BATCH:
            raise ValueError(f"train_steps_host: batch {batch} > {PERSIST_MAX_BATCH}; use train_step_host_batch per step")
        if self.world_size > 1 and self.comm is None:
            raise RuntimeError("train_steps_host at world_size > 1 needs enable_fused_ddp()")
        if slots is None:                         # largest power of two whose ring stays under 64 MB (1024 at batch 64):
            slots = 16                            # each launch covers slots/2 steps, a launch boundary costs ~30 us
            while slots < 4096 and 2 * slots * batch * 785 <= (64 << 20):
                slots *= 2
        feed = self._feeds.get((batch, slots))
        if feed is None:
            feed = self._feeds[(batch, slots)] = ops.HostFeed(batch, slots)
        st0 = self.step_count if philox_step0 is None else philox_step0
        self.step_count += nsteps
        mk = dict(masks=masks) if masks is not None else dict(philox=(self.seed, st0))
        ops.train_steps_host_fed(feed, self.params, self.ws, nsteps, host_pixels=host_pixels, host_labels=host_labels,
                                 comm=self.comm, lr=self.lr, loss=self.loss, host_loss_steps=host_loss_steps, **mk)

    def check_feed(self) -> None:
        for feed in self._feeds.values():
            if feed.error():
                raise RuntimeError("host-fed training loop: the kernel waited ~3 s for a batch that never arrived")

    def train_step_host_batch(self, x_u8: torch.Tensor, y_u8: torch.Tensor, idx_identity: torch.Tensor,
                              mask: Optional[torch.Tensor] = None) -> None:
        """Same step for a batch that was just copied from host memory (u8 [B,784] + u8 [B]):
        the staged batch is used as a B-sample dataset indexed by `idx_identity` = arange(B)."""
        class _Staged:
            images, labels = x_u8, y_u8
        b = x_u8.shape[0]
        from ._lib import PERSIST_MAX_BATCH
        if b <= PERSIST_MAX_BATCH:
            self.train_steps(_Staged, idx_identity, b, 1, masks=None if mask is None else mask[None])
        else:
            self.train_step(_Staged, idx_identity, mask=mask)

    def eval_step(self, resident, idx: torch.Tensor) -> None:
        """Validation forward (mp:472-481): eval mode, loss only, accumulates loss[1]."""
        ops.step_forward(self.params, self.ws, idx.numel(), images=resident.images, labels=resident.labels,
                         idx=idx, loss=self.loss)

    def zero_epoch_loss(self) -> None:
        self.loss.zero_()

    def epoch_loss(self) -> float:
        """Host read (one sync): sum over batches of mean_loss_b / B_b (mp:465, 481)."""
        return float(self.loss[1].item())

    def last_batch_loss(self) -> float:
        return float(self.loss[0].item())

    # ------------------------------------------------------------------ CUDA graph of K steps
    def capture_steps(self, resident, idx_batches: Sequence[torch.Tensor],
                      masks: Optional[Sequence[torch.Tensor]] = None, philox_step0: int = 0,
                      persistent: bool = False) -> torch.cuda.CUDAGraph:
        """Capture the given steps, unrolled, into one CUDA graph (the batch-64 regime is
        launch-bound: 3 kernels of a few microseconds per step).  The graph reads `idx_batches`'
        storage at replay time, so refilling that storage replays the same steps on new indices."""
        if self.world_size > 1:
            import torch.distributed as dist
            dist.all_reduce(self.grads, group=self.group)          # warm NCCL outside the capture
            self.grads.zero_()
        torch.cuda.synchronize(self.device)
        g = torch.cuda.CUDAGraph()
        saved = self.step_count
        with torch.cuda.graph(g):
            for t, idx in enumerate(idx_batches):
                if persistent:
                    self.train_steps(resident, idx, idx.numel(), 1,
                                     masks=None if masks is None else masks[t][None], philox_step0=philox_step0 + t)
                else:
                    self.train_step(resident, idx, mask=None if masks is None else masks[t],
                                    philox_step=philox_step0 + t)
        self.step_count = saved + len(idx_batches)
        return g
