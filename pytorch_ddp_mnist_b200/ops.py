"""Tensor-level wrappers over the C ABI: argument checks + pointer plumbing, nothing else.

Every function launches on torch's current CUDA stream and returns without synchronising.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import _lib
from ._lib import CLASSES, HID, IN, N_PARAMS, check, cur_stream, lib, ptr

__all__ = ["decode_lut", "gather_decode", "dropout_mask_philox", "Workspace", "step_forward",
           "step_backward", "train_steps_persistent", "train_steps_fused_ddp", "FusedComm", "cross_entropy",
           "sgd_step"]


def _need(t: Optional[torch.Tensor], name: str, dtype, shape=None, optional=False):
    if t is None:
        if optional:
            return
        raise ValueError(f"{name} is required")
    if not t.is_cuda:
        raise ValueError(f"{name} must be a CUDA tensor (this library has no CPU path)")
    if t.dtype != dtype:
        raise TypeError(f"{name} must be {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError(f"{name} must be contiguous")
    if shape is not None and tuple(t.shape) != tuple(shape):
        raise ValueError(f"{name} must have shape {tuple(shape)}, got {tuple(t.shape)}")


def decode_lut() -> torch.Tensor:
    """256-entry fp32 decode table, computed by the library on the host."""
    out = torch.empty(256, dtype=torch.float32)
    check(lib.ddpm_decode_lut_host(out.data_ptr()))
    return out


def gather_decode(images: torch.Tensor, labels: torch.Tensor, idx: torch.Tensor,
                  out_x: Optional[torch.Tensor] = None, out_y: Optional[torch.Tensor] = None,
                  status: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """(x f32[B,784], y u8[B]) for sample indices idx i64[B] of the resident u8 dataset."""
    n = images.shape[0]
    _need(images, "images", torch.uint8)
    if images.numel() != n * IN:
        raise ValueError("images must hold 784 bytes per sample")
    _need(labels, "labels", torch.uint8, (n,))
    _need(idx, "idx", torch.int64)
    b = idx.numel()
    if out_x is None:
        out_x = torch.empty((b, IN), dtype=torch.float32, device=images.device)
    if out_y is None:
        out_y = torch.empty((b,), dtype=torch.uint8, device=images.device)
    _need(out_x, "out_x", torch.float32, (b, IN))
    _need(out_y, "out_y", torch.uint8, (b,))
    _need(status, "status", torch.int32, optional=True)
    check(lib.ddpm_gather_decode(ptr(images), ptr(labels), n, ptr(idx), b, ptr(out_x), ptr(out_y),
                                 ptr(status), cur_stream()))
    return out_x, out_y


def dropout_mask_philox(batch: int, seed: int, step: int, device, out: Optional[torch.Tensor] = None
                        ) -> torch.Tensor:
    if out is None:
        out = torch.empty((batch, HID), dtype=torch.uint8, device=device)
    _need(out, "out", torch.uint8, (batch, HID))
    check(lib.ddpm_dropout_mask_philox(ptr(out), batch, seed, step, cur_stream()))
    return out


class Workspace:
    """Caller-owned scratch for the fused step (see ddpm_workspace_bytes in the header)."""

    def __init__(self, max_batch: int, device):
        self.max_batch = int(max_batch)
        nbytes = lib.ddpm_workspace_bytes(self.max_batch)
        if nbytes == 0:
            raise ValueError("max_batch must be positive")
        # torch's caching allocator hands out >= 512-byte aligned blocks
        self.buf = torch.empty(nbytes, dtype=torch.uint8, device=device)
        if self.buf.data_ptr() % 256:
            raise RuntimeError("workspace is not 256-byte aligned")
        check(lib.ddpm_workspace_init(ptr(self.buf), self.max_batch, cur_stream()))

    @property
    def nbytes(self) -> int:
        return self.buf.numel()

    def peek(self, batch: int, which: str) -> torch.Tensor:
        """Test aid: one of the per-layer arrays the three-kernel forward / backward left in the workspace
        (`h1d`, `h2`, `dz2`, `dz1`: [batch,128]; `dz3`: [batch,16])."""
        names = ("h1d", "h2", "dz2", "dz1", "dz3")
        w = names.index(which)
        out = torch.empty((batch, 128 if w < 4 else 16), dtype=torch.float32, device=self.buf.device)
        check(lib.ddpm_workspace_peek(ptr(self.buf), batch, w, ptr(out), cur_stream()))
        return out

    def error(self) -> int:
        """Non-zero once a wait inside the persistent kernel gave up (~3 s): see ddpm_workspace_error.  Blocking."""
        import ctypes as C
        e = C.c_int(0)
        check(lib.ddpm_workspace_error(ptr(self.buf), C.byref(e)))
        return e.value


def step_forward(params: torch.Tensor, ws: Workspace, batch: int, *,
                 images: Optional[torch.Tensor] = None, labels: Optional[torch.Tensor] = None,
                 idx: Optional[torch.Tensor] = None, x: Optional[torch.Tensor] = None,
                 y: Optional[torch.Tensor] = None, mask: Optional[torch.Tensor] = None,
                 philox: Optional[Tuple[int, int]] = None, prep_backward: bool = False,
                 logits: Optional[torch.Tensor] = None, loss: Optional[torch.Tensor] = None,
                 tf32: bool = False) -> None:
    _need(params, "params", torch.float32, (N_PARAMS,))
    if batch > ws.max_batch:
        raise ValueError(f"batch {batch} exceeds workspace max_batch {ws.max_batch}")
    n = 0
    if x is not None:
        _need(x, "x", torch.float32, (batch, IN))
    else:
        _need(images, "images", torch.uint8)
        _need(idx, "idx", torch.int64)
        if idx.numel() < batch:
            raise ValueError("idx shorter than batch")
        n = images.shape[0]
        _need(labels, "labels", torch.uint8, (n,), optional=True)
    _need(y, "y", torch.uint8, optional=True)
    _need(mask, "mask", torch.uint8, (batch, HID), optional=True)
    _need(logits, "logits", torch.float32, (batch, CLASSES), optional=True)
    _need(loss, "loss", torch.float32, (2,), optional=True)
    flags = ((_lib.FWD_PREP_BACKWARD if prep_backward else 0) | (_lib.FWD_PHILOX_MASK if philox else 0)
             | (_lib.FWD_TF32 if tf32 else 0))
    seed, step = philox if philox else (0, 0)
    check(lib.ddpm_step_forward(ptr(params), ptr(images), ptr(labels), n, ptr(idx), ptr(x), ptr(y),
                                ptr(mask), seed, step, batch, flags, ptr(logits), ptr(loss),
                                ptr(ws.buf), cur_stream()))


def step_backward(params: torch.Tensor, ws: Workspace, batch: int, *,
                  images: Optional[torch.Tensor] = None, idx: Optional[torch.Tensor] = None,
                  x: Optional[torch.Tensor] = None, dlogits: Optional[torch.Tensor] = None,
                  grads: Optional[torch.Tensor] = None, grad_scale: float = 1.0, lr: float = 0.0,
                  tf32: bool = False, part: Optional[str] = None) -> None:
    _need(params, "params", torch.float32, (N_PARAMS,))
    _need(grads, "grads", torch.float32, (N_PARAMS,), optional=True)
    _need(dlogits, "dlogits", torch.float32, (batch, CLASSES), optional=True)
    n = 0
    if x is not None:
        _need(x, "x", torch.float32, (batch, IN))
    else:
        _need(images, "images", torch.uint8)
        _need(idx, "idx", torch.int64)
        n = images.shape[0]
    check(lib.ddpm_step_backward(ptr(params), ptr(images), n, ptr(idx), ptr(x), ptr(dlogits), batch,
                                 grad_scale, lr, (_lib.BWD_TF32 if tf32 else 0) | {None: 0, "small": _lib.BWD_PART_SMALL, "w1": _lib.BWD_PART_W1}[part],
                                 ptr(grads), ptr(ws.buf), cur_stream()))


def train_steps_persistent(params: torch.Tensor, ws: Workspace, batch: int, nsteps: int, *,
                           images: torch.Tensor, labels: torch.Tensor, idx: torch.Tensor,
                           masks: Optional[torch.Tensor] = None, philox: Optional[Tuple[int, int]] = None,
                           grads: Optional[torch.Tensor] = None, grad_scale: float = 1.0, lr: float = 0.0,
                           loss: Optional[torch.Tensor] = None, loss_steps: Optional[torch.Tensor] = None) -> None:
    """`nsteps` whole optimisation steps in one cooperative launch (batch <= 512).
    idx: i64[nsteps*batch] consecutive batches; masks: u8[nsteps,batch,128] or None."""
    _need(params, "params", torch.float32, (N_PARAMS,))
    if batch > ws.max_batch:
        raise ValueError(f"batch {batch} exceeds workspace max_batch {ws.max_batch}")
    _need(images, "images", torch.uint8)
    n = images.shape[0]
    _need(labels, "labels", torch.uint8, (n,))
    _need(idx, "idx", torch.int64)
    if idx.numel() < nsteps * batch:
        raise ValueError("idx shorter than nsteps*batch")
    _need(masks, "masks", torch.uint8, (nsteps, batch, HID), optional=True)
    _need(grads, "grads", torch.float32, (N_PARAMS,), optional=True)
    _need(loss, "loss", torch.float32, (2,), optional=True)
    _need(loss_steps, "loss_steps", torch.float32, optional=True)
    if loss_steps is not None and loss_steps.numel() < nsteps:
        raise ValueError("loss_steps shorter than nsteps")
    seed, step0 = philox if philox else (0, 0)
    flags = _lib.FWD_PHILOX_MASK if philox else 0
    check(lib.ddpm_train_steps_persistent(ptr(params), ptr(grads), ptr(images), ptr(labels), n, ptr(idx),
                                          ptr(masks), seed, step0, batch, nsteps, flags, grad_scale, lr,
                                          ptr(loss), ptr(loss_steps), ptr(ws.buf), cur_stream()))


def exchange_only(comm: "FusedComm", nsteps: int, out: Optional[torch.Tensor] = None) -> None:
    """The fused step's gradient exchange alone, `nsteps` times (measurement / test aid, collective)."""
    _need(out, "out", torch.float32, (N_PARAMS,), optional=True)
    check(lib.ddpm_comm_exchange_only(comm._h, int(nsteps), ptr(out), cur_stream()))


def launch_floor(nlaunches: int = 1) -> None:
    """`nlaunches` empty cooperative launches with the step kernel's launch shape (measurement aid)."""
    check(lib.ddpm_persist_launch_floor(int(nlaunches), cur_stream()))


class FusedComm:
    """Communicator of the fused data-parallel step (ddpm_comm_*): a symmetric device buffer per
    rank, mapped into every peer process with CUDA IPC.  Handles travel over torch.distributed."""

    def __init__(self, rank: int, world: int, group=None):
        """Collective over `group`.  Every failure point is followed by an agreement step (all-reduce of a success
        flag), so either every rank gets a connected communicator or every rank raises `DDPMError` - no rank is
        left waiting in a collective for a peer that gave up."""
        import ctypes as C
        import torch.distributed as dist
        self.rank, self.world = rank, world
        self._h = C.c_void_p()
        dev = torch.device("cuda", torch.cuda.current_device())

        def agree(ok: bool, what: str, err: str = ""):
            flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
            if not bool(flag.item()):
                self.close()
                raise _lib.DDPMError(f"fused data-parallel communicator: {what} failed on "
                                     f"{'this rank: ' + err if not ok else 'another rank'}")

        err = ""
        ok = lib.ddpm_comm_create(rank, world, C.byref(self._h)) == 0
        if not ok:
            err, self._h = _lib.last_error(), C.c_void_p()
        agree(ok, "ddpm_comm_create", err)
        mine = (C.c_ubyte * 64)()
        ok = lib.ddpm_comm_ipc_handle(self._h, mine) == 0
        agree(ok, "ddpm_comm_ipc_handle", "" if ok else _lib.last_error())
        gathered = [None] * world
        dist.all_gather_object(gathered, bytes(mine), group=group)
        blob = (C.c_ubyte * (64 * world)).from_buffer_copy(b"".join(gathered))
        ok = lib.ddpm_comm_connect(self._h, blob) == 0
        agree(ok, "ddpm_comm_connect (CUDA IPC mapping of a peer's buffer)", "" if ok else _lib.last_error())
        dist.barrier(group=group)                         # everyone is mapped before anyone launches

    @staticmethod
    def usable(world: int, group=None) -> Tuple[bool, str]:
        """Collective: can this job use the in-kernel exchange?  It needs every rank on ONE host (CUDA IPC), at most
        8 ranks (one NVLink domain) and peer access between every pair of the job's GPUs.  All ranks return the same
        answer; the string says why not."""
        import socket
        import torch.distributed as dist
        if world > 8:
            return False, f"world size {world} > 8"
        me = (socket.gethostname(), torch.cuda.current_device())
        everyone = [None] * world
        dist.all_gather_object(everyone, me, group=group)
        if len({h for h, _ in everyone}) != 1:
            return False, "ranks span more than one host"
        mine = me[1]
        ok = all(d == mine or torch.cuda.can_device_access_peer(mine, d) for _, d in everyone)
        if len({d for _, d in everyone}) != world:
            ok = False                                    # two ranks on one GPU
        flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=torch.device("cuda", mine))
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        return (True, "") if bool(flag.item()) else (False, "no peer access between some pair of GPUs")

    def error(self) -> int:
        import ctypes as C
        e = C.c_int(0)
        check(lib.ddpm_comm_error(self._h, C.byref(e)))
        return e.value

    def close(self) -> None:
        if self._h:
            lib.ddpm_comm_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def train_steps_fused_ddp(comm: FusedComm, params: torch.Tensor, ws: Workspace, batch: int, nsteps: int, *,
                          images: torch.Tensor, labels: torch.Tensor, idx: torch.Tensor,
                          idx_all: Optional[torch.Tensor] = None,
                          masks: Optional[torch.Tensor] = None, philox: Optional[Tuple[int, int]] = None,
                          lr: float, loss: Optional[torch.Tensor] = None,
                          loss_steps: Optional[torch.Tensor] = None) -> None:
    """`nsteps` data-parallel optimisation steps in one launch per rank; the gradient average and
    the update happen inside the kernel over NVLink peer memory.  `idx_all` i64[world, >= nsteps*batch]
    (every rank's batches) enables the dz1 exchange for fc1 when batch <= 64."""
    _need(params, "params", torch.float32, (N_PARAMS,))
    if batch > ws.max_batch:
        raise ValueError(f"batch {batch} exceeds workspace max_batch {ws.max_batch}")
    _need(images, "images", torch.uint8)
    n = images.shape[0]
    _need(labels, "labels", torch.uint8, (n,))
    _need(idx, "idx", torch.int64)
    if idx.numel() < nsteps * batch:
        raise ValueError("idx shorter than nsteps*batch")
    _need(masks, "masks", torch.uint8, (nsteps, batch, HID), optional=True)
    _need(loss, "loss", torch.float32, (2,), optional=True)
    _need(loss_steps, "loss_steps", torch.float32, optional=True)
    seed, step0 = philox if philox else (0, 0)
    flags = _lib.FWD_PHILOX_MASK if philox else 0
    stride = 0
    if idx_all is not None:
        _need(idx_all, "idx_all", torch.int64)
        if idx_all.dim() != 2 or idx_all.shape[0] != comm.world or idx_all.shape[1] < nsteps * batch:
            raise ValueError("idx_all must be i64[world, >= nsteps*batch]")
        stride = idx_all.shape[1]
    check(lib.ddpm_train_steps_fused_ddp(comm._h, ptr(params), ptr(images), ptr(labels), n, ptr(idx), ptr(idx_all),
                                         stride, ptr(masks), seed, step0, batch, nsteps, flags, lr, ptr(loss),
                                         ptr(loss_steps), ptr(ws.buf), cur_stream()))


class HostFeed:
    """Ring of batch slots in device memory fed from pinned host memory while the persistent kernel
    runs (ddpm_feed_*): the reference's `for x, y in train_loader: x.to(device) ... loss.item()` loop
    (mp:452-466) as one call."""

    def __init__(self, batch: int, slots: int = 256):
        import ctypes as C
        self.batch, self.slots = int(batch), int(slots)
        self._h = C.c_void_p()
        check(lib.ddpm_feed_create(self.batch, self.slots, C.byref(self._h)))

    def error(self) -> int:
        import ctypes as C
        e = C.c_int(0)
        check(lib.ddpm_feed_error(self._h, C.byref(e)))
        return e.value

    def close(self) -> None:
        if self._h:
            lib.ddpm_feed_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _need_pinned(t: Optional[torch.Tensor], name: str, dtype, optional=False):
    if t is None:
        if optional:
            return
        raise ValueError(f"{name} is required")
    if t.is_cuda or not t.is_pinned():
        raise ValueError(f"{name} must be a pinned host tensor")
    if t.dtype != dtype:
        raise TypeError(f"{name} must be {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError(f"{name} must be contiguous")


def train_steps_host_fed(feed: HostFeed, params: torch.Tensor, ws: Workspace, nsteps: int, *,
                         host_pixels: torch.Tensor, host_labels: torch.Tensor,
                         comm: Optional[FusedComm] = None,
                         masks: Optional[torch.Tensor] = None, philox: Optional[Tuple[int, int]] = None,
                         lr: float, loss: Optional[torch.Tensor] = None,
                         host_loss_steps: Optional[torch.Tensor] = None) -> None:
    """`nsteps` optimisation steps on batches that live in PINNED HOST memory (`host_pixels`
    u8[nsteps, batch, 784], `host_labels` u8[nsteps, batch]): every step's batch is copied to the
    device while the kernel runs and every step's loss is written to `host_loss_steps` (pinned
    f32[nsteps]).  The host buffers must stay alive until the current stream has passed this call."""
    _need(params, "params", torch.float32, (N_PARAMS,))
    b = feed.batch
    if b > ws.max_batch:
        raise ValueError(f"batch {b} exceeds workspace max_batch {ws.max_batch}")
    _need_pinned(host_pixels, "host_pixels", torch.uint8)
    _need_pinned(host_labels, "host_labels", torch.uint8)
    if host_pixels.numel() < nsteps * b * IN or host_labels.numel() < nsteps * b:
        raise ValueError("host buffers shorter than nsteps*batch samples")
    _need(masks, "masks", torch.uint8, (nsteps, b, HID), optional=True)
    _need(loss, "loss", torch.float32, (2,), optional=True)
    _need_pinned(host_loss_steps, "host_loss_steps", torch.float32, optional=True)
    if host_loss_steps is not None and host_loss_steps.numel() < nsteps:
        raise ValueError("host_loss_steps shorter than nsteps")
    seed, step0 = philox if philox else (0, 0)
    flags = _lib.FWD_PHILOX_MASK if philox else 0
    check(lib.ddpm_train_steps_host_fed(feed._h, comm._h if comm is not None else None, ptr(params),
                                        ptr(host_pixels), ptr(host_labels), ptr(masks), seed, step0, nsteps, flags, lr,
                                        ptr(loss), ptr(host_loss_steps), ptr(ws.buf), cur_stream()))


def cross_entropy(logits: torch.Tensor, y: torch.Tensor, loss: Optional[torch.Tensor] = None,
                  dlogits: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    _need(logits, "logits", torch.float32)
    b = logits.shape[0]
    if tuple(logits.shape) != (b, CLASSES):
        raise ValueError("logits must be [B,10]")
    _need(y, "y", torch.uint8, (b,))
    if loss is None:
        loss = torch.empty(1, dtype=torch.float32, device=logits.device)
    _need(dlogits, "dlogits", torch.float32, (b, CLASSES), optional=True)
    check(lib.ddpm_cross_entropy(ptr(logits), ptr(y), b, ptr(loss), ptr(dlogits), cur_stream()))
    return loss, dlogits


def sgd_step(params: torch.Tensor, grads: torch.Tensor, lr: float, grad_scale: float = 1.0) -> None:
    _need(params, "params", torch.float32)
    _need(grads, "grads", torch.float32, tuple(params.shape))
    check(lib.ddpm_sgd_step(ptr(params), ptr(grads), params.numel(), lr, grad_scale, cur_stream()))
