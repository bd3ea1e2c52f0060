"""Determinisation shared by the golden-fixture generator and the tests (TEST INFRASTRUCTURE).

The reference seeds nothing but its sampler, so dropout masks are injected: training forward
number `t` uses `bernoulli(0.8)` drawn from `torch.Generator().manual_seed(MASK_SEED0 + t)`.
"""
import numpy as np
import torch
from torch import nn

MASK_SEED0 = 10_000
MODEL_SEED = 0


def mask_for_step(t: int, batch: int, width: int = 128, keep: float = 0.8) -> torch.Tensor:
    g = torch.Generator().manual_seed(MASK_SEED0 + t)
    return torch.bernoulli(torch.full((batch, width), keep), generator=g)


class MaskFeed:
    """Replaces nn.Dropout.forward with the mask-injecting form; forward value and gradient are
    `x * mask * 1.25`, identical to ATen's native_dropout for a given mask.  The step counter
    advances once per TRAINING forward."""

    def __init__(self):
        self.step = 0
        self.masks = []

    def install(self):
        feed = self

        def forward(mod, x):
            if not mod.training:
                return x
            mask = mask_for_step(feed.step, x.shape[0], x.shape[1], 1.0 - mod.p)
            feed.masks.append(mask.numpy().astype(np.uint8))
            feed.step += 1
            return x * (mask * (1.0 / (1.0 - mod.p)))
        self._orig = nn.Dropout.forward
        nn.Dropout.forward = forward
        return self

    def remove(self):
        nn.Dropout.forward = self._orig
