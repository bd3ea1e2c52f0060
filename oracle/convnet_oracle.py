"""CPU oracle for the BASELINE.json "ConvNet" rows — TEST INFRASTRUCTURE ONLY.

PARITY UNPINNED.  The model these rows describe (conv 1->32 3x3, conv 32->64 3x3, maxpool 2,
dropout .25, fc 9216->128, dropout .5, fc 128->10, log_softmax + nll_loss, Adadelta) does NOT exist
anywhere under /root/reference: there is no reference code, test or golden vector to
pin this file against.  It is a CPU fp32 restatement of the layer list BASELINE.json gives, written
with torch.nn.functional + autograd, and every result checked against it must be reported as
"oracle = restatement of BASELINE.json, not reference".

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import numpy as np
import torch
import torch.nn.functional as F

PARAM_KEYS = ("conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias",
              "fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias")
PARAM_SHAPES = ((32, 1, 3, 3), (32,), (64, 32, 3, 3), (64,), (128, 9216), (128,), (10, 128), (10,))
N_PARAMS = 1_199_882
KEEP1, KEEP2 = 0.75, 0.5                 # Dropout(0.25) after the pool, Dropout(0.5) after fc1
RHO, EPS, LR = 0.9, 1e-6, 1.0            # Adadelta defaults of the BASELINE.json recipe (row J5)


def param_offsets() -> Tuple[int, ...]:
    offs, o = [], 0
    for s in PARAM_SHAPES:
        offs.append(o)
        o += int(np.prod(s))
    assert o == N_PARAMS
    return tuple(offs)


def init_params(seed: int = 0) -> Dict[str, np.ndarray]:
    """torch's default initialisation of the same layers (kaiming-uniform weights, uniform biases)."""
    g = torch.Generator().manual_seed(seed)
    out = {}
    for k, s in zip(PARAM_KEYS, PARAM_SHAPES):
        fan_in = int(np.prod(s[1:])) if len(s) > 1 else None
        if fan_in is None:                              # bias: fan_in of its weight
            w = out[k.replace("bias", "weight")]
            fan_in = int(np.prod(w.shape[1:]))
        bound = 1.0 / np.sqrt(fan_in)
        out[k] = ((torch.rand(s, generator=g) * 2 - 1) * bound).numpy().astype(np.float32)
    return out


def flatten_params(p: Dict[str, np.ndarray]) -> np.ndarray:
    return np.concatenate([np.asarray(p[k], np.float32).reshape(-1) for k in PARAM_KEYS])


def unflatten_params(flat: np.ndarray) -> Dict[str, np.ndarray]:
    out, o = {}, 0
    for k, s in zip(PARAM_KEYS, PARAM_SHAPES):
        n = int(np.prod(s))
        out[k] = np.asarray(flat[o:o + n], np.float32).reshape(s)
        o += n
    return out


def forward_backward(p: Dict[str, np.ndarray], x: np.ndarray, y: np.ndarray,
                     mask1: Optional[np.ndarray] = None, mask2: Optional[np.ndarray] = None,
                     want_grads: bool = True, dtype=np.float32):
    """One step's forward (+ backward).  x f32[B,784] (decoded), y u8[B]; mask1 u8[B,9216] /
    mask2 u8[B,128] are the dropout keep masks (None = eval mode).  Returns (loss, log_probs[B,10],
    grads dict or None, intermediates dict).  dtype=np.float64 evaluates the same fp32 inputs in double
    precision: the yardstick that tells summation-order noise (of the CPU's fp32 run as much as of the
    GPU's) from an error."""
    t = {k: torch.tensor(np.asarray(v, np.float32).astype(dtype), requires_grad=want_grads) for k, v in p.items()}
    xb = torch.from_numpy(np.asarray(x, np.float32).astype(dtype)).reshape(-1, 1, 28, 28)
    a1 = F.relu(F.conv2d(xb, t["conv1.weight"], t["conv1.bias"]))              # J1
    a2 = F.relu(F.conv2d(a1, t["conv2.weight"], t["conv2.bias"]))              # J2
    p2 = torch.flatten(F.max_pool2d(a2, 2), 1)
    p2d = p2 if mask1 is None else p2 * torch.from_numpy(mask1.astype(dtype)) * float(np.float32(1.0 / KEEP1))
    h = F.relu(F.linear(p2d, t["fc1.weight"], t["fc1.bias"]))                  # J3
    hd = h if mask2 is None else h * torch.from_numpy(mask2.astype(dtype)) * float(np.float32(1.0 / KEEP2))
    z = F.linear(hd, t["fc2.weight"], t["fc2.bias"])
    lp = F.log_softmax(z, dim=1)
    loss = F.nll_loss(lp, torch.from_numpy(np.asarray(y)).long())
    grads = None
    if want_grads:
        loss.backward()                                                        # J4
        grads = {k: t[k].grad.numpy().copy() for k in PARAM_KEYS}
    inter = {"a1": a1.detach().numpy(), "a2": a2.detach().numpy(), "p2": p2.detach().numpy(), "h": h.detach().numpy()}
    return float(loss.detach()), lp.detach().numpy(), grads, inter


def adadelta_step(p: np.ndarray, g: np.ndarray, sq: np.ndarray, acc: np.ndarray,
                  lr: float = LR, rho: float = RHO, eps: float = EPS):
    """torch.optim.Adadelta's single-tensor rule (torch/optim/adadelta.py:292-302), fp32, on flat arrays.
    Returns (p, sq, acc) updated."""
    f = np.float32
    p, g, sq, acc = (np.asarray(a, f) for a in (p, g, sq, acc))
    sq = (sq * f(rho) + (f(1 - rho) * g) * g).astype(f)            # addcmul_: self + value * t1 * t2
    std = np.sqrt(sq + f(eps), dtype=f)
    delta = (np.sqrt(acc + f(eps), dtype=f) / std * g).astype(f)
    acc = (acc * f(rho) + (f(1 - rho) * delta) * delta).astype(f)
    p = (p - f(lr) * delta).astype(f)
    return p, sq, acc


def pool_near_ties(a2: np.ndarray, rel: float = 1e-5) -> int:
    """Number of 2x2 pooling windows of a2 [B,64,24,24] whose maximum is positive and whose two largest values
    agree to within `rel` (relative): there an fp32 evaluation in another summation order may pick the other
    element as the argmax and route the window's gradient to a neighbouring position."""
    b, c = a2.shape[:2]
    win = a2.reshape(b, c, 12, 2, 12, 2).transpose(0, 1, 2, 4, 3, 5).reshape(b, c, 12, 12, 4)
    srt = np.sort(win, -1)
    top, second = srt[..., 3], srt[..., 2]
    return int(((top > 0) & ((top - second) <= rel * np.abs(top))).sum())
