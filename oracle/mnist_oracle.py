"""CPU oracle for the data-parallel MNIST training step — TEST INFRASTRUCTURE, NOT A PRODUCT PATH.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference` legs
may import this module, and only as the checker (or the timed CPU baseline).  Nothing under
`pytorch_ddp_mnist_b200/` imports it; the product path fails loudly when the CUDA library is
missing and has no CPU fallback.

What it is: a numpy/fp32 restatement of every arithmetic stage of the reference's step
(`/root/reference/mnist_pnetcdf_cpu_mp.py`, cited per function as `mp:LINE`), written from the
formulas, not from torch modules, so that it can run anywhere numpy runs.

Parity status: PINNED.  `tests/test_oracle.py` checks every function here against fixtures that
`oracle/gen_golden.py` produced by running the reference's own code in the build container
(tests/golden/*.npz): decode table bit-exact, sampler shards bit-exact, step-0 activations / loss /
gradients, the SGD update bit-exact, a 100-step ws=1 trajectory, a 20-step ws=2 gloo trajectory.
The one piece taken from a third-party library rather than restated is the permutation itself:
`DistributedSampler` is *defined* as torch's CPU mt19937 `randperm` (torch >= 1.6 pinned by the
reference's requirements.txt:1; 2.11.0 here), so `shard_indices` calls `torch.randperm`.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

F32 = np.float32
PARAM_KEYS = ("0.weight", "0.bias", "3.weight", "3.bias", "5.weight")     # state_dict order, mp:412-422
PARAM_SHAPES = ((128, 784), (128,), (128, 128), (128,), (10, 128))
N_PARAMS = 118_272
MEAN, STD = F32(0.1307), F32(0.3081)                                       # mp:376
KEEP = 0.8                                                                 # Dropout(0.2), mp:417
SCALE = F32(1.0 / KEEP)                                                    # 1.25 exactly
LR = F32(0.01)                                                             # mp:443


# ------------------------------------------------------------------ R1: sampler  (mp:390-394,449)
def shard_indices(n: int, world_size: int, rank: int, epoch: int = 0, seed: int = 42) -> np.ndarray:
    """`DistributedSampler(shuffle=True, seed, drop_last=False)` after `set_epoch(epoch)`:
    perm = randperm(n, mt19937(seed+epoch)); pad by wrapping to ceil(n/ws)*ws; take rank::ws.
    (torch/utils/data/distributed.py:102-141.)"""
    import torch
    g = torch.Generator()
    g.manual_seed(seed + epoch)
    perm = torch.randperm(n, generator=g).numpy().astype(np.int64)
    per = -(-n // world_size)
    total = per * world_size
    pad = total - n
    if pad > 0:
        reps = -(-pad // n)
        perm = np.concatenate([perm, np.tile(perm, reps)[:pad]])
    return perm[rank:total:world_size]


def batches(shard: np.ndarray, batch: int) -> List[np.ndarray]:
    """`DataLoader(batch_size, drop_last=False)` over the shard: consecutive chunks, short tail
    (mp:396-401)."""
    return [shard[s:s + batch] for s in range(0, len(shard), batch)]


# ------------------------------------------------------------------ R2/R3: gather + decode
def decode_lut() -> np.ndarray:
    """fp32 value of each byte after `ToTensor` then `Normalize((0.1307,),(0.3081,))` (mp:374-377):
    true IEEE fp32 operations in this order: v/255, -0.1307f, /0.3081f."""
    v = np.arange(256, dtype=F32)
    return ((v / F32(255.0)) - MEAN) / STD


def gather_decode(images_u8: np.ndarray, labels_u8: np.ndarray, idx: np.ndarray
                  ) -> Tuple[np.ndarray, np.ndarray]:
    """`MNISTNetCDF.__getitem__` over a batch of indices + default_collate + `x.view(B,-1)`
    (mp:39-49, 458): returns (x f32[B,784], y u8[B])."""
    flat = images_u8.reshape(images_u8.shape[0], -1)
    return decode_lut()[flat[idx]], labels_u8[idx].astype(np.uint8)


# ------------------------------------------------------------------ R6-R8: forward  (mp:412-422)
def forward(p: Dict[str, np.ndarray], x: np.ndarray, mask: Optional[np.ndarray]) -> Dict[str, np.ndarray]:
    """mask=None is eval mode (dropout is the identity, mp:470); otherwise mask is {0,1}[B,128]."""
    a = {}
    a["z1"] = x @ p["0.weight"].T + p["0.bias"]                 # Linear(784,128)
    a["h1"] = np.maximum(a["z1"], F32(0))                       # ReLU
    a["h1d"] = a["h1"] if mask is None else a["h1"] * (mask.astype(F32) * SCALE)   # Dropout(0.2)
    a["z2"] = a["h1d"] @ p["3.weight"].T + p["3.bias"]          # Linear(128,128)
    a["h2"] = np.maximum(a["z2"], F32(0))                       # ReLU
    a["z3"] = a["h2"] @ p["5.weight"].T                         # Linear(128,10,bias=False)
    return a


# ------------------------------------------------------------------ R9: CrossEntropyLoss (mp:444,461)
def cross_entropy(z: np.ndarray, y: np.ndarray) -> Tuple[np.float32, np.ndarray]:
    """Mean-reduced log_softmax + NLL on raw logits; y may be uint8 (the loader's dtype).
    Returns (loss, dL/dz = (softmax(z) - onehot(y)) / B)."""
    b = z.shape[0]
    m = z.max(axis=1, keepdims=True)
    e = np.exp(z - m)
    s = e.sum(axis=1, keepdims=True)
    logp = (z - m) - np.log(s)
    yi = y.astype(np.int64)
    loss = F32(-(logp[np.arange(b), yi].astype(np.float64).sum()) / b)
    dz = e / s
    dz[np.arange(b), yi] -= F32(1)
    return loss, (dz / F32(b)).astype(F32)


# ------------------------------------------------------------------ R10: backward (mp:462)
def backward(p: Dict[str, np.ndarray], x: np.ndarray, a: Dict[str, np.ndarray],
             mask: Optional[np.ndarray], dz3: np.ndarray) -> Dict[str, np.ndarray]:
    g = {}
    g["5.weight"] = dz3.T @ a["h2"]
    dh2 = dz3 @ p["5.weight"]
    dz2 = dh2 * (a["z2"] > 0)
    g["3.weight"] = dz2.T @ a["h1d"]
    g["3.bias"] = dz2.sum(axis=0)
    dh1d = dz2 @ p["3.weight"]
    dh1 = dh1d if mask is None else dh1d * (mask.astype(F32) * SCALE)
    dz1 = dh1 * (a["z1"] > 0)
    g["0.weight"] = dz1.T @ x
    g["0.bias"] = dz1.sum(axis=0)
    return {k: v.astype(F32) for k, v in g.items()}


# ------------------------------------------------------------------ R11: DDP gradient mean (mp:439)
def ddp_mean(grads_per_rank: Sequence[Dict[str, np.ndarray]]) -> Dict[str, np.ndarray]:
    """c10d reducer semantics: each rank's gradient is multiplied by 1/ws while it is copied into
    the single 473,088-byte bucket, then the buckets are summed."""
    ws = len(grads_per_rank)
    inv = F32(1.0 / ws)
    out = {}
    for k in PARAM_KEYS:
        acc = grads_per_rank[0][k] * inv
        for r in range(1, ws):
            acc = acc + grads_per_rank[r][k] * inv
        out[k] = acc.astype(F32)
    return out


# ------------------------------------------------------------------ R12: SGD (mp:443,463)
def sgd_step(p: Dict[str, np.ndarray], g: Dict[str, np.ndarray], lr: np.float32 = LR) -> Dict[str, np.ndarray]:
    """`param.add_(grad, alpha=-lr)`: ATen evaluates it as ONE fused multiply-add per element
    (pinned bit-exactly by tests/golden/traj_ws1.npz params0/grads0/params1), i.e.
    fp32(p + (-lr)*g) with a single rounding; float64 holds the fp32 product exactly."""
    nlr = np.float64(F32(-lr))
    return {k: (p[k].astype(np.float64) + nlr * g[k].astype(np.float64)).astype(F32) for k in PARAM_KEYS}


# ------------------------------------------------------------------ whole step / loops
def train_step(p, x, y, mask, peers: Optional[Sequence[Tuple[np.ndarray, np.ndarray, np.ndarray]]] = None,
               lr: np.float32 = LR):
    """One optimisation step on this rank's batch.  `peers` = other ranks' (x, y, mask) for the
    same step when emulating DDP in one process.  Returns (new_params, loss, grads, acts)."""
    a = forward(p, x, mask)
    loss, dz3 = cross_entropy(a["z3"], y)
    g = backward(p, x, a, mask, dz3)
    if peers:
        gs = [g]
        for (xp, yp, mp_) in peers:
            ap = forward(p, xp, mp_)
            _, dzp = cross_entropy(ap["z3"], yp)
            gs.append(backward(p, xp, ap, mp_, dzp))
        g = ddp_mean(gs)
    return sgd_step(p, g, lr), loss, g, a


def eval_loss(p, x, y) -> np.float32:
    """Validation forward (mp:469-480): eval mode, no dropout."""
    return cross_entropy(forward(p, x, None)["z3"], y)[0]


def epoch_loss_sum(batch_losses: Sequence[float], batch_sizes: Sequence[int]) -> float:
    """The reference's bookkeeping, kept as is: `epoch_loss += batch_loss / x.shape[0]`
    (mp:465, 481) — an already-mean loss divided by the batch size again."""
    return float(sum(float(l) / b for l, b in zip(batch_losses, batch_sizes)))


def flatten_params(p: Dict[str, np.ndarray]) -> np.ndarray:
    """The flat 118,272-element layout shared with the CUDA library: W1,b1,W2,b2,W3 in
    `model.parameters()` order, each row-major [out,in]."""
    return np.concatenate([p[k].reshape(-1) for k in PARAM_KEYS]).astype(F32)


def unflatten_params(flat: np.ndarray) -> Dict[str, np.ndarray]:
    out, pos = {}, 0
    for k, shp in zip(PARAM_KEYS, PARAM_SHAPES):
        n = int(np.prod(shp))
        out[k] = flat[pos:pos + n].reshape(shp).astype(F32)
        pos += n
    return out


def unpack_masks(packed: np.ndarray) -> np.ndarray:
    """Fixture masks are np.packbits(..., axis=-1) of {0,1}[steps,B,128]."""
    return np.unpackbits(packed, axis=-1).astype(np.uint8)
