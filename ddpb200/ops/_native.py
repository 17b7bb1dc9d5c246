"""ctypes binding of libddpb200.so (the in-tree sm_100a kernel library).

The reference is pure Python on ATen CPU ops; every op it uses has a
hand-written kernel here instead, reached through a plain C ABI (include/ddpb200.h) so
that non-Python callers can use the same entry points.

There is deliberately NO silent fallback: on a CUDA device an op either runs its kernel
from this library or raises.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import torch

_LIB = None
_LIB_PATH = Path(os.environ.get("DDPB_LIB") or (Path(__file__).resolve().parent.parent / "_lib" / "libddpb200.so"))

c_void = C.c_void_p
c_int = C.c_int
c_ll = C.c_longlong
c_float = C.c_float

# name -> argtypes (restype is int unless listed in _RESTYPES)
_SIGNATURES = {
    "ddpb_abi_version": [],
    "ddpb_sm_count": [],
    "ddpb_conv_fprop": [c_void, c_void, c_void, c_void, c_void, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                        c_void, c_ll, c_void, c_void],
    "ddpb_conv_split_sync_words": [],
    "ddpb_conv_fuse_split": [c_int],
    "ddpb_conv_fprop_scratch_elems": [c_int, c_int, c_int, c_int, c_int, c_int, c_int],
    "ddpb_conv_bn_relu_pool_fwd": [c_void] * 9 + [c_int] * 7 + [c_float, c_void, c_ll, c_void, c_void],
    "ddpb_conv_force": [c_int, c_int, c_int],
    "ddpb_conv_debug_timing": [c_void],
    "ddpb_conv_debug_mode": [c_int],
    "ddpb_conv_tail_from_tmem": [c_int],
    "ddpb_conv_plan": [c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void],
    "ddpb_conv_plan_ex": [c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void],
    "ddpb_conv_wgrad_splits": [c_int, c_int, c_int, c_int, c_int, c_int],
    "ddpb_conv_wgrad_scratch_elems": [c_int, c_int, c_int, c_int, c_int, c_int],
    "ddpb_conv_wgrad_scratch_elems_ld": [c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int],
    "ddpb_conv_wgrad": [c_void, c_void, c_void, c_void, c_ll, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                        c_void],
    "ddpb_channel_stats": [c_void, c_void, c_ll, c_int, c_void],
    "ddpb_bn_relu_pool_fwd": [c_void, c_void, c_void, c_void, c_void, c_void, c_int, c_int, c_int, c_int, c_int,
                              c_float, c_void],
    "ddpb_bn_relu_pool_bwd": [c_void, c_void, c_void, c_void, c_void, c_void, c_void, c_void, c_void, c_void,
                              c_void, c_ll, c_int, c_int, c_int, c_int, c_int, c_float, c_void],
    "ddpb_bn_bwd_partials_elems": [c_int, c_int, c_int, c_int, c_int],
    "ddpb_bn_bwd_set_fused": [c_int],
    "ddpb_stamp": [c_void, c_void],
    "ddpb_publish_scalar": [c_void, c_void, c_void, c_void],
    "ddpb_set_pdl": [c_int],
    "ddpb_get_pdl": [],
    "ddpb_fc_ce_fwd": [c_void, c_void, c_void, c_void, c_void, c_void, c_void, c_void, c_void, c_void, c_int,
                       c_int, c_int, c_void],
    "ddpb_fc_bwd": [c_void, c_void, c_void, c_void, c_void, c_void, c_int, c_int, c_int, c_void, c_void, c_void,
                    c_void, c_void],
    "ddpb_sgd_step": [c_void, c_void, c_void, c_void, c_ll, c_float, c_float, c_float, c_float, c_void],
    "ddpb_conv_dgrad": [c_void, c_void, c_void, c_int, c_int, c_int, c_int, c_int, c_int, c_void, c_ll, c_void, c_void],
    "ddpb_pack_weights": [c_int, c_void, c_void, c_void, c_void, c_void, c_void, c_void, c_void, c_void],
    "ddpb_first_layer_operand_f32": [c_void, c_void, c_int, c_int, c_int, c_void, c_void, c_int, c_void],
    "ddpb_first_layer_operand_u8": [c_void, c_void, c_void, c_int, c_int, c_int, c_void, c_void, c_int, c_void,
                                    c_void, c_int, c_void],
    "ddpb_draw_augmentation": [c_void, c_int, c_int, C.c_ulonglong, c_void, c_void],
    "ddpb_synthetic_batch": [c_void, c_void, c_int, c_int, c_int, c_int, C.c_ulonglong, c_void, c_void],
}
_RESTYPES = {
    "ddpb_last_error": C.c_char_p,
    "ddpb_conv_wgrad_scratch_elems": c_ll,
    "ddpb_conv_wgrad_scratch_elems_ld": c_ll,
    "ddpb_conv_fprop_scratch_elems": c_ll,
}
# entry points added by later translation units; bound if present
_OPTIONAL: dict[str, list] = {}


def register_optional(name: str, argtypes: list, restype=c_int) -> None:
    _OPTIONAL[name] = argtypes
    if restype is not c_int:
        _RESTYPES[name] = restype
    if _LIB is not None and hasattr(_LIB, name):
        fn = getattr(_LIB, name)
        fn.argtypes = argtypes
        fn.restype = restype


def lib_path() -> Path:
    return _LIB_PATH


def available() -> bool:
    return _LIB_PATH.exists()


def load(build_if_missing: bool = True) -> C.CDLL:
    """Loads (building first if needed and possible) the kernel library."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if not _LIB_PATH.exists() or os.environ.get("DDPB_REBUILD") == "1":
        if not build_if_missing:
            raise RuntimeError(f"{_LIB_PATH} is missing; run `python -m ddpb200.build`")
        from .. import build as _build

        _build.build()
    lib = C.CDLL(str(_LIB_PATH))
    lib.ddpb_last_error.restype = C.c_char_p
    lib.ddpb_last_error.argtypes = []
    for name, args in {**_SIGNATURES, **_OPTIONAL}.items():
        if not hasattr(lib, name):
            if name in _SIGNATURES:
                raise RuntimeError(f"{_LIB_PATH} lacks {name}; stale build? run `python -m ddpb200.build --force`")
            continue
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = _RESTYPES.get(name, c_int)
    _LIB = lib
    return lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().ddpb_last_error()
        raise RuntimeError(f"{what} failed (code {rc}): {msg.decode() if msg else '?'}")


def ptr(t: torch.Tensor | None) -> int | None:
    """Device pointer of a tensor (None -> NULL)."""
    if t is None:
        return None
    return t.data_ptr()


def stream_ptr(stream: torch.cuda.Stream | None = None) -> int:
    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream


def require_cuda(*tensors: torch.Tensor) -> None:
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError("ddpb200 native ops need CUDA tensors; use ddpb200.ops.reference on CPU")
