"""Drop-in `create_model()`, loss and optimizer: the reference's PyTorch object protocol
 on top of the CUDA step kernels.

With these three swapped in, the reference's loop body runs textually unchanged (mp:455-465):

    x = x.view(x.shape[0], -1); optimizer.zero_grad(); y_hat = model(x)
    batch_loss = loss(y_hat, y); batch_loss.backward(); optimizer.step(); batch_loss.item()

`model` keeps nn.Sequential's surface (`.to()`, `.train()/.eval()`, `.parameters()` = W1,b1,W2,b2,W3,
`state_dict()` keys `0.weight 0.bias 3.weight 3.bias 5.weight`, survives `DistributedDataParallel`)
but its five parameters are views into ONE flat 118,272-float buffer and its forward/backward are
the fused kernels.  CUDA only: a CPU tensor raises (there is no CPU fallback).
"""
from __future__ import annotations

from typing import Callable, Optional

import torch
from torch import nn

from . import ops
from ._lib import N_PARAMS
from .engine import PARAM_OFFSETS, PARAM_SHAPES


class _MLPFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, module, *params):
        if not x.is_cuda:
            raise RuntimeError("pytorch_ddp_mnist_b200 model got a CPU tensor: this library has no CPU path")
        x = x.contiguous().float()
        b = x.shape[0]
        flat, ws = module._flat, module._workspace(b)
        logits = torch.empty((b, 10), dtype=torch.float32, device=x.device)
        kw = {}
        if module.training:
            if module._mask_provider is not None:
                kw["mask"] = module._mask_provider(module._train_forwards, b).to(x.device, torch.uint8).contiguous()
            else:
                kw["philox"] = (module._seed, module._train_forwards)
            module._train_forwards += 1
        need_bwd = any(ctx.needs_input_grad)          # grad mode is off inside Function.forward
        ops.step_forward(flat, ws, b, x=x, prep_backward=need_bwd, logits=logits, **kw)
        # the activations backward needs live in the module's ONE workspace, not in ctx: stamp this forward so that a
        # backward which comes after another forward (validation inside the step, micro-batches, two losses) fails
        # loudly instead of returning the gradients of the wrong batch
        module._fwd_generation += 1
        ctx.module, ctx.batch, ctx.generation = module, b, module._fwd_generation
        ctx.save_for_backward(x)
        return logits

    @staticmethod
    def backward(ctx, dlogits):
        (x,) = ctx.saved_tensors
        m = ctx.module
        if m._fwd_generation != ctx.generation:
            raise RuntimeError("pytorch_ddp_mnist_b200: backward() of a forward pass whose activations were overwritten by "
                               "a later forward of the same model (the model keeps one workspace: call backward() "
                               "before the next model(x), as the reference's loop does, mp:460-462)")
        g = torch.empty(N_PARAMS, dtype=torch.float32, device=x.device)
        ops.step_backward(m._flat, m._workspace(ctx.batch), ctx.batch, x=x,
                          dlogits=dlogits.contiguous().float(), grads=g)
        m._last_grads = g
        views = []
        for shp, off in zip(PARAM_SHAPES, PARAM_OFFSETS):
            n = 1
            for s in shp:
                n *= s
            views.append(g[off:off + n].view(shp))
        return (None, None, *views)


class B200MLP(nn.Sequential):
    """`create_model()`'s network (mp:412-422) with flat storage and fused CUDA forward/backward."""

    def __init__(self):
        super().__init__(nn.Linear(28 * 28, 128), nn.ReLU(), nn.Dropout(0.2), nn.Linear(128, 128), nn.ReLU(),
                         nn.Linear(128, 10, bias=False))
        self._flat: Optional[torch.Tensor] = None
        self._ws = None
        self._mask_provider: Optional[Callable[[int, int], torch.Tensor]] = None
        self._train_forwards = 0
        self._seed = 42
        self._last_grads = None
        self._fwd_generation = 0

    # ---- flat storage ------------------------------------------------------------------
    def _param_list(self):
        return [self[0].weight, self[0].bias, self[3].weight, self[3].bias, self[5].weight]

    def _flatten(self) -> None:
        ps = self._param_list()
        dev = ps[0].device
        if dev.type != "cuda":
            self._flat = None
            return
        flat = torch.empty(N_PARAMS, dtype=torch.float32, device=dev)
        for p, shp, off in zip(ps, PARAM_SHAPES, PARAM_OFFSETS):
            v = flat[off:off + p.numel()].view(shp)
            v.copy_(p.data)
            p.data = v
        self._flat = flat
        self._ws = None

    def _is_flat(self) -> bool:
        if self._flat is None:
            return False
        base = self._flat.data_ptr()
        return all(p.data_ptr() == base + 4 * off and p.is_contiguous()
                   for p, off in zip(self._param_list(), PARAM_OFFSETS))

    def _apply(self, fn, *a, **k):
        out = super()._apply(fn, *a, **k)
        self._flatten()
        return out

    def _workspace(self, batch: int):
        if self._ws is None or self._ws.max_batch < batch or self._ws.buf.device != self._flat.device:
            self._ws = ops.Workspace(max(batch, 128), self._flat.device)
        return self._ws

    # ---- knobs ---------------------------------------------------------------------------
    def set_mask_provider(self, fn: Optional[Callable[[int, int], torch.Tensor]]) -> None:
        """Parity mode: `fn(train_forward_number, batch) -> {0,1}[batch,128]` supplies the dropout
        mask (the oracle's); None = in-kernel Philox stream (perf mode)."""
        self._mask_provider = fn

    def set_dropout_seed(self, seed: int) -> None:
        self._seed = int(seed)

    # ---- forward -------------------------------------------------------------------------
    def forward(self, x):
        if not self._is_flat():
            self._flatten()
            if self._flat is None:
                raise RuntimeError("move the model to a CUDA device first: this library has no CPU path")
        return _MLPFunction.apply(x, self, *self._param_list())


def create_model() -> nn.Module:
    """Same call as the reference's `create_model()` (mp:412-422); same initial values under the
    same `torch.manual_seed` (the submodules are constructed in the same order)."""
    return B200MLP()


class _CEFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, target):
        if not logits.is_cuda:
            raise RuntimeError("CrossEntropyLoss got a CPU tensor: this library has no CPU path")
        y = target if target.dtype == torch.uint8 else target.to(torch.uint8)
        loss = torch.empty(1, dtype=torch.float32, device=logits.device)
        dl = torch.empty_like(logits, dtype=torch.float32) if ctx.needs_input_grad[0] else None
        ops.cross_entropy(logits.contiguous().float(), y.contiguous(), loss=loss, dlogits=dl)
        if dl is not None:
            ctx.save_for_backward(dl)
        return loss[0]

    @staticmethod
    def backward(ctx, grad_out):
        (dl,) = ctx.saved_tensors
        return dl * grad_out, None


class CrossEntropyLoss(nn.Module):
    """`nn.CrossEntropyLoss()` as the reference uses it (mp:444, 461): class-index targets (uint8
    from the loader, or int64), mean reduction, raw logits [B,10]."""

    def forward(self, input, target):
        return _CEFunction.apply(input, target)


class SGD(torch.optim.Optimizer):
    """`optim.SGD(params, lr)` with no momentum / weight decay (all the reference uses, mp:443):
    `p = fma(-lr, g, p)`, one kernel over the flat buffer when parameters and gradients are flat."""

    def __init__(self, params, lr: float = 0.01):
        if lr < 0:
            raise ValueError(f"Invalid learning rate: {lr}")
        super().__init__(params, dict(lr=lr))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group in self.param_groups:
            ps = [p for p in group["params"] if p.grad is not None]
            if not ps:
                continue
            lr = group["lr"]
            if len(ps) == 5 and sum(p.numel() for p in ps) == N_PARAMS and self._flat_pair(ps):
                n0 = ps[0]
                flat = torch.as_strided(n0.data, (N_PARAMS,), (1,), n0.data.storage_offset())
                gflat = torch.as_strided(n0.grad, (N_PARAMS,), (1,), n0.grad.storage_offset())
                ops.sgd_step(flat, gflat, lr)
            else:
                for p in ps:
                    ops.sgd_step(p.data.view(-1), p.grad.contiguous().view(-1), lr)
        return loss

    @staticmethod
    def _flat_pair(ps) -> bool:
        pb, gb, pos = ps[0].data_ptr(), ps[0].grad.data_ptr(), 0
        for p in ps:
            if not (p.is_contiguous() and p.grad.is_contiguous() and p.grad.dtype == torch.float32):
                return False
            if p.data_ptr() != pb + 4 * pos or p.grad.data_ptr() != gb + 4 * pos:
                return False
            pos += p.numel()
        g0 = ps[0].grad
        if g0.untyped_storage().nbytes() < 4 * (g0.storage_offset() + N_PARAMS):
            return False
        return pb % 16 == 0 and gb % 16 == 0
