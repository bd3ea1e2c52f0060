"""ctypes binding of libddpmnist_b200.so (include/ddpmnist_b200.h).

There is no CPU fallback anywhere in this package: if the shared library is missing or a symbol
is absent, importing this module raises, and every op raises `DDPMError` with the library's own
message when a call fails.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

_HERE = os.path.dirname(os.path.abspath(__file__))
# DDPM_LIB: another build of the same library (kernel experiments: scripts build variants next to the default one)
LIB_PATH = os.environ.get("DDPM_LIB") or os.path.join(_HERE, "libddpmnist_b200.so")

IN, HID, CLASSES, N_PARAMS = 784, 128, 10, 118272
OFF_W1, OFF_B1, OFF_W2, OFF_B2, OFF_W3 = 0, 100352, 100480, 116864, 116992
FWD_PREP_BACKWARD, FWD_PHILOX_MASK, FWD_TF32 = 1, 2, 4
BWD_TF32, BWD_PART_SMALL, BWD_PART_W1 = 1, 2, 4
CN_BWD_TRAIN, CN_BWD_TF32, CN_BWD_FWD_WAS_TF32 = 1, 2, 4
STATIC_SCRATCH_MAX_BATCH = 128
PERSIST_MAX_BATCH = 512
CN_N_PARAMS, CN_MAX_BATCH = 1199882, 8192
ABI_VERSION = 6


class DDPMError(RuntimeError):
    pass


def _load() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or pytorch_ddp_mnist_b200/csrc/build.sh (needs nvcc; there is no CPU fallback)")
    return C.CDLL(LIB_PATH)


lib = _load()

_vp, _i64, _i32, _u64, _f32, _u32 = C.c_void_p, C.c_int64, C.c_int, C.c_uint64, C.c_float, C.c_uint

# name -> (restype, argtypes); every symbol declared in include/ddpmnist_b200.h
SIGNATURES = {
    "ddpm_last_error": (C.c_char_p, []),
    "ddpm_abi_version": (_i32, []),
    "ddpm_decode_lut_host": (_i32, [_vp]),
    "ddpm_gather_decode": (_i32, [_vp, _vp, _i64, _vp, _i32, _vp, _vp, _vp, _vp]),
    "ddpm_dropout_mask_philox": (_i32, [_vp, _i32, _u64, _u64, _vp]),
    "ddpm_workspace_bytes": (C.c_size_t, [_i32]),
    "ddpm_workspace_init": (_i32, [_vp, _i32, _vp]),
    "ddpm_workspace_forward_bytes": (C.c_size_t, [_i32]),
    "ddpm_workspace_peek": (_i32, [_vp, _i32, _i32, _vp, _vp]),
    "ddpm_step_forward": (_i32, [_vp, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _u64, _u64, _i32, _u32,
                                 _vp, _vp, _vp, _vp]),
    "ddpm_step_forward_static_scratch": (_i32, [_vp, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _u64, _u64, _i32,
                                                _u32, _vp, _vp, _vp]),
    "ddpm_train_steps_persistent": (_i32, [_vp, _vp, _vp, _vp, _i64, _vp, _vp, _u64, _u64, _i32, _i32, _u32,
                                           _f32, _f32, _vp, _vp, _vp, _vp]),
    "ddpm_persist_launch_floor": (_i32, [_i32, _vp]),
    "ddpm_workspace_error": (_i32, [_vp, C.POINTER(_i32)]),
    "ddpm_comm_exchange_only": (_i32, [_vp, _i32, _vp, _vp]),
    "ddpm_comm_create": (_i32, [_i32, _i32, C.POINTER(_vp)]),
    "ddpm_comm_ipc_handle": (_i32, [_vp, _vp]),
    "ddpm_comm_connect": (_i32, [_vp, _vp]),
    "ddpm_comm_error": (_i32, [_vp, C.POINTER(_i32)]),
    "ddpm_comm_destroy": (None, [_vp]),
    "ddpm_train_steps_fused_ddp": (_i32, [_vp, _vp, _vp, _vp, _i64, _vp, _vp, _i64, _vp, _u64, _u64, _i32, _i32, _u32,
                                          _f32, _vp, _vp, _vp, _vp]),
    "ddpm_feed_create": (_i32, [_i32, _i32, C.POINTER(_vp)]),
    "ddpm_feed_error": (_i32, [_vp, C.POINTER(_i32)]),
    "ddpm_feed_destroy": (None, [_vp]),
    "ddpm_train_steps_host_fed": (_i32, [_vp, _vp, _vp, _vp, _vp, _vp, _u64, _u64, _i32, _u32, _f32, _vp, _vp, _vp, _vp]),
    "ddpm_cn_workspace_bytes": (C.c_size_t, [_i32]),
    "ddpm_cn_workspace_init": (_i32, [_vp, _i32, _vp]),
    "ddpm_cn_step_forward": (_i32, [_vp, _vp, _vp, _i64, _vp, _vp, _vp, _u64, _u64, _i32, _u32, _vp, _vp, _vp, _vp]),
    "ddpm_cn_step_backward": (_i32, [_vp, _vp, _i64, _vp, _i32, _u32, _f32, _vp, _vp, _vp]),
    "ddpm_adadelta_step": (_i32, [_vp, _vp, _vp, _vp, _i64, _f32, C.c_double, C.c_double, _f32, _vp]),
    "ddpm_step_backward": (_i32, [_vp, _vp, _i64, _vp, _vp, _vp, _i32, _f32, _f32, _u32, _vp, _vp, _vp]),
    "ddpm_cross_entropy": (_i32, [_vp, _vp, _i32, _vp, _vp, _vp]),
    "ddpm_sgd_step": (_i32, [_vp, _vp, _i64, _f32, _f32, _vp]),
    "ddpm_cdf5_open": (_i32, [C.c_char_p, C.POINTER(_vp)]),
    "ddpm_cdf5_close": (None, [_vp]),
    "ddpm_cdf5_inq_var": (_i32, [_vp, C.c_char_p, C.POINTER(_i32), C.POINTER(_i64), C.POINTER(_i32),
                                 C.POINTER(_i32), C.POINTER(_i64)]),
    "ddpm_cdf5_read_rows": (_i32, [_vp, C.c_char_p, _i64, _i64, _vp]),
}

for _name, (_res, _args) in SIGNATURES.items():
    try:
        _fn = getattr(lib, _name)
    except AttributeError as e:                                    # pragma: no cover
        raise ImportError(f"{LIB_PATH} does not export {_name}; rebuild it") from e
    _fn.restype = _res
    _fn.argtypes = _args

if lib.ddpm_abi_version() != ABI_VERSION:                           # pragma: no cover
    raise ImportError(f"{LIB_PATH} has ABI {lib.ddpm_abi_version()}, bindings expect {ABI_VERSION}; rebuild")


def last_error() -> str:
    return (lib.ddpm_last_error() or b"").decode("utf-8", "replace")


def check(rc: int) -> None:
    if rc != 0:
        raise DDPMError(last_error())


def ptr(t) -> Optional[int]:
    """Device/host address of a torch tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()


def cur_stream() -> int:
    import torch
    return torch.cuda.current_stream().cuda_stream
