"""BASELINE.json's "ConvNet"  behind the same boundary as the MLP.

NOT IN THE REFERENCE: /root/reference defines only the 784-128-128-10 MLP.  This module
drives the `ddpm_cn_*` / `ddpm_adadelta_step` entry points (include/ddpmnist_b200.h) for the layer list
BASELINE.json names — conv 1->32, conv 32->64, maxpool 2, dropout .25, fc 9216->128, dropout .5,
fc 128->10, log_softmax + nll_loss, Adadelta — over the same resident u8 dataset and sampler shards.
Checked against oracle/convnet_oracle.py (a restatement: parity unpinned).  No CPU fallback.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch

from . import _lib
from ._lib import CN_N_PARAMS, check, cur_stream, lib, ptr
from .ops import _need

PARAM_KEYS = ("conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias",
              "fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias")
PARAM_SHAPES = ((32, 1, 3, 3), (32,), (64, 32, 3, 3), (64,), (128, 9216), (128,), (10, 128), (10,))
FEAT, FC = 9216, 128


def init_params_like_torch() -> Dict[str, torch.Tensor]:
    """torch's default initialisation of the layers BASELINE.json names (current RNG state), under the
    state_dict keys of the usual `Net` module."""
    import torch.nn as nn
    layers = {"conv1": nn.Conv2d(1, 32, 3, 1), "conv2": nn.Conv2d(32, 64, 3, 1), "fc1": nn.Linear(9216, 128),
              "fc2": nn.Linear(128, 10)}
    return {f"{n}.{k}": v.detach().clone() for n, m in layers.items() for k, v in m.state_dict().items()}


class ConvNetWorkspace:
    """Caller-owned scratch of the ConvNet step (ddpm_cn_workspace_bytes: ~0.66 MB per image + 11 MB)."""

    def __init__(self, max_batch: int, device):
        self.max_batch = int(max_batch)
        nbytes = lib.ddpm_cn_workspace_bytes(self.max_batch)
        if nbytes == 0:
            raise ValueError(f"max_batch must be in [1, {_lib.CN_MAX_BATCH}]")
        self.buf = torch.empty(nbytes, dtype=torch.uint8, device=device)
        if self.buf.data_ptr() % 256:
            raise RuntimeError("workspace is not 256-byte aligned")
        check(lib.ddpm_cn_workspace_init(ptr(self.buf), self.max_batch, cur_stream()))


def step_forward(params: torch.Tensor, ws: ConvNetWorkspace, *, images: torch.Tensor, labels: torch.Tensor,
                 idx: torch.Tensor, mask1: Optional[torch.Tensor] = None, mask2: Optional[torch.Tensor] = None,
                 philox: Optional[Tuple[int, int]] = None, prep_backward: bool = False,
                 logprobs: Optional[torch.Tensor] = None, loss: Optional[torch.Tensor] = None,
                 tf32: bool = False) -> None:
    """conv1 .. nll_loss for the samples `idx` of the resident dataset.  Dropout: `mask1` u8[B,9216] +
    `mask2` u8[B,128] (injected), or `philox=(seed, step)`, or neither = eval mode.  `tf32`: conv2 as an
    implicit GEMM on the tcgen05 tensor cores (TF32 operands, fp32 accumulate; ~1e-3 relative)."""
    _need(params, "params", torch.float32, (CN_N_PARAMS,))
    _need(images, "images", torch.uint8)
    n = images.shape[0]
    _need(labels, "labels", torch.uint8, (n,))
    _need(idx, "idx", torch.int64)
    b = idx.numel()
    if b > ws.max_batch:
        raise ValueError(f"batch {b} exceeds workspace max_batch {ws.max_batch}")
    _need(mask1, "mask1", torch.uint8, (b, FEAT), optional=True)
    _need(mask2, "mask2", torch.uint8, (b, FC), optional=True)
    _need(logprobs, "logprobs", torch.float32, (b, 10), optional=True)
    _need(loss, "loss", torch.float32, (2,), optional=True)
    flags = ((_lib.FWD_PREP_BACKWARD if prep_backward else 0) | (_lib.FWD_PHILOX_MASK if philox else 0)
             | (_lib.FWD_TF32 if tf32 else 0))
    seed, step = philox if philox else (0, 0)
    check(lib.ddpm_cn_step_forward(ptr(params), ptr(images), ptr(labels), n, ptr(idx), ptr(mask1), ptr(mask2),
                                   seed, step, b, flags, ptr(logprobs), ptr(loss), ptr(ws.buf), cur_stream()))


def step_backward(params: torch.Tensor, ws: ConvNetWorkspace, grads: torch.Tensor, *, images: torch.Tensor,
                  idx: torch.Tensor, train: bool = True, grad_scale: float = 1.0, tf32: bool = False,
                  forward_was_tf32: bool = False) -> None:
    """Gradient bucket (f32[1199882], scaled by `grad_scale`) of the step whose forward just ran with
    `prep_backward=True` on `ws`.  `tf32`: conv2's gradients on the tensor cores; `forward_was_tf32`: that
    forward also ran with tf32, so its TF32 operand copies in the workspace are reused."""
    _need(params, "params", torch.float32, (CN_N_PARAMS,))
    _need(grads, "grads", torch.float32, (CN_N_PARAMS,))
    _need(images, "images", torch.uint8)
    _need(idx, "idx", torch.int64)
    check(lib.ddpm_cn_step_backward(ptr(params), ptr(images), images.shape[0], ptr(idx), idx.numel(),
                                    (_lib.CN_BWD_TRAIN if train else 0) | (_lib.CN_BWD_TF32 if tf32 else 0)
                                    | (_lib.CN_BWD_FWD_WAS_TF32 if (tf32 and forward_was_tf32) else 0), grad_scale, ptr(grads), ptr(ws.buf), cur_stream()))


def adadelta_step(params: torch.Tensor, grads: torch.Tensor, square_avg: torch.Tensor, acc_delta: torch.Tensor,
                  lr: float = 1.0, rho: float = 0.9, eps: float = 1e-6, grad_scale: float = 1.0) -> None:
    """torch.optim.Adadelta's update, fused, on flat fp32 arrays of equal length."""
    n = params.numel()
    for t, name in ((params, "params"), (grads, "grads"), (square_avg, "square_avg"), (acc_delta, "acc_delta")):
        _need(t, name, torch.float32, (n,))
    check(lib.ddpm_adadelta_step(ptr(params), ptr(grads), ptr(square_avg), ptr(acc_delta), n, lr, rho, eps,
                                 grad_scale, cur_stream()))


class ConvNetEngine:
    """Flat-parameter training engine for the ConvNet: forward, backward into one bucket, (all-reduce), Adadelta."""

    def __init__(self, device, max_batch: int, lr: float = 1.0, rho: float = 0.9, eps: float = 1e-6,
                 world_size: int = 1, rank: int = 0, group=None, seed: int = 42, tf32: bool = False):
        self.device = torch.device(device)
        self.tf32 = bool(tf32)
        self.lr, self.rho, self.eps = float(lr), float(rho), float(eps)
        self.world_size, self.rank, self.group = int(world_size), int(rank), group
        self.seed = int(seed) + 0x9E3779B97F4A7C15 * self.rank & 0xFFFFFFFFFFFFFFFF
        z = lambda: torch.zeros(CN_N_PARAMS, dtype=torch.float32, device=self.device)
        self.params, self.grads, self.square_avg, self.acc_delta = z(), z(), z(), z()
        self.loss = torch.zeros(2, dtype=torch.float32, device=self.device)
        self.ws = ConvNetWorkspace(max_batch, self.device)
        self.step_count = 0

    def views(self) -> Dict[str, torch.Tensor]:
        out, o = {}, 0
        for k, shp in zip(PARAM_KEYS, PARAM_SHAPES):
            n = 1
            for s in shp:
                n *= s
            out[k] = self.params[o:o + n].view(shp)
            o += n
        return out

    def load_state_dict(self, sd) -> None:
        v = self.views()
        for k in PARAM_KEYS:
            v[k].copy_(torch.as_tensor(sd[k]).to(torch.float32))

    def state_dict(self) -> Dict[str, torch.Tensor]:
        return {k: t.detach().cpu().clone() for k, t in self.views().items()}

    def broadcast_params(self, src: int = 0) -> None:
        if self.world_size > 1:
            import torch.distributed as dist
            dist.broadcast(self.params, src=src, group=self.group)

    def train_step(self, resident, idx: torch.Tensor, mask1: Optional[torch.Tensor] = None,
                   mask2: Optional[torch.Tensor] = None) -> None:
        if mask1 is None:
            kw = dict(philox=(self.seed, self.step_count))
        else:
            kw = dict(mask1=mask1, mask2=mask2)
        self.step_count += 1
        step_forward(self.params, self.ws, images=resident.images, labels=resident.labels, idx=idx, prep_backward=True,
                     loss=self.loss, tf32=self.tf32, **kw)
        step_backward(self.params, self.ws, self.grads, images=resident.images, idx=idx, train=True,
                      grad_scale=1.0 / self.world_size, tf32=self.tf32, forward_was_tf32=self.tf32)
        if self.world_size > 1:
            import torch.distributed as dist
            dist.all_reduce(self.grads, op=dist.ReduceOp.SUM, group=self.group)
        adadelta_step(self.params, self.grads, self.square_avg, self.acc_delta, self.lr, self.rho, self.eps)

    def capture_steps(self, resident, idx_batches) -> "torch.cuda.CUDAGraph":
        """The given steps, unrolled, in ONE CUDA graph (the batch-64 step is 13 dependent launches of a few microseconds
        each: launch gaps were ~40 % of it).  The graph reads the storage of `idx_batches` at replay time - refill it and
        replay to train on new samples.  Dropout uses the Philox counters step_count .. step_count + len - 1 captured
        here on every replay (benchmark use; a training loop that needs fresh counters re-captures per epoch)."""
        if self.world_size > 1:
            import torch.distributed as dist
            dist.all_reduce(self.grads, group=self.group)          # warm NCCL outside the capture
            self.grads.zero_()
        torch.cuda.synchronize(self.device)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for idx in idx_batches:
                self.train_step(resident, idx)
        return g

    def eval_step(self, resident, idx: torch.Tensor, logprobs: Optional[torch.Tensor] = None) -> None:
        step_forward(self.params, self.ws, images=resident.images, labels=resident.labels, idx=idx, loss=self.loss,
                     logprobs=logprobs)

    def scale_lr(self, gamma: float) -> None:
        """`StepLR(step_size=1, gamma)` of the recipe BASELINE.json describes: call once per epoch."""
        self.lr *= float(gamma)

    def accuracy(self, resident, idx: torch.Tensor) -> float:
        """Fraction of the samples `idx` whose arg-max class equals the label (eval mode; the test loop's metric)."""
        lp = torch.empty(idx.numel(), 10, dtype=torch.float32, device=self.device)
        self.eval_step(resident, idx, logprobs=lp)
        return float((lp.argmax(dim=1) == resident.labels[idx].to(torch.int64)).float().mean().item())

    def last_batch_loss(self) -> float:
        return float(self.loss[0].item())
