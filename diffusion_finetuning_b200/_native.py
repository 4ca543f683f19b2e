"""ctypes binding of libb200lora.so (the C-ABI declared in include/b200lora.h).

The library is the ONLY compute path for CUDA tensors: if it is missing or a call fails this
module raises — there is no eager/PyTorch fallback behind it.
"""
import ctypes
import os
import threading

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200lora.so")

F32, F16, BF16 = 0, 1, 2
IMPL_AUTO, IMPL_SIMT, IMPL_TC = 0, 1, 2
STATIC_SCRATCH_BYTES = 4 << 20      # B200LORA_STATIC_SCRATCH_BYTES (include/b200lora.h)
W_NK, W_KN = 0, 1

_DTYPE_CODE = {torch.float32: F32, torch.float16: F16, torch.bfloat16: BF16}

_lib = None
_lock = threading.Lock()

_vp, _i64, _i32, _f32 = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_float

# name -> (restype, argtypes); mirrors include/b200lora.h one to one
SIGNATURES = {
    "b200lora_version": (_i32, []),
    "b200lora_last_error": (ctypes.c_char_p, []),
    "b200lora_tc_supported": (_i32, []),
    "b200lora_watchdog_code": (ctypes.c_uint, []),
    "b200lora_peek_cuda_error": (_i32, []),
    "b200lora_debug_timeline": (None, [_vp, _i32]),
    "b200lora_lora_gemm": (_i32, [_vp, _i64, _vp, _i64, _i32, _vp, _i32, _vp, _i64, _i64, _vp, _vp,
                                  _i64, _i64, _i32, _vp, _f32, _vp, _i64, _vp, _i64, _i32, _i32,
                                  _i32, _i32, _i32, _vp]),
    "b200lora_pack_elems": (_i64, [_i32, _i32]),
    "b200lora_pack_lora": (_i32, [_vp, _vp, _i32, _vp, _i32, _vp, _i32, _i32, _i32, _i32, _vp]),
    "b200lora_pack_lora_fused": (_i32, [_vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _vp, _i32, _i32, _i32, _vp]),
    "b200lora_pack_lora_cast": (_i32, [_vp, _vp, _i32, _vp, _i32, _vp, _i32, _i32, _i32, _i32, _vp, _vp, _i64, _vp]),
    "b200lora_pack_lora_fused_cast": (_i32, [_vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _vp, _i32, _i32, _i32, _vp, _vp, _i64, _vp]),
    "b200lora_linear_fwd": (_i32, [_vp, _vp, _vp, _i32, _vp, _vp, _i32, _vp, _f32, _vp, _vp, _i64,
                                   _i32, _i32, _i32, _i32, _i32, _vp]),
    "b200lora_linear_bwd_dx": (_i32, [_vp, _vp, _i32, _vp, _vp, _i32, _vp, _f32, _vp, _vp, _i64,
                                      _i32, _i32, _i32, _i32, _i32, _vp]),
    "b200lora_linear_bwd_dx_f32out": (_i32, [_vp, _vp, _vp, _f32, _vp, _vp, _i64, _i32, _i32, _i32, _i32, _vp]),
    "b200lora_linear_bwd_dx_fused_f32out": (_i32, [_vp, _vp, _i32, _vp, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _i32, _vp]),
    "b200lora_linear_fwd_geglu": (_i32, [_vp, _vp, _vp, _i32, _f32, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _i32, _vp]),
    "b200lora_geglu_bwd": (_i32, [_vp, _vp, _vp, _i64, _i32, _i32, _vp]),
    "b200lora_lora_grads": (_i32, [_vp, _vp, _vp, _vp, _f32, _vp, _vp, _i64, _i32, _i32, _i32,
                                   _i32, _i32, _vp, _i64, _vp]),
    "b200lora_lora_grads_workspace": (_i64, [_i64, _i32, _i32, _i32]),
    "b200lora_lora_grads_static": (_i32, [_vp, _vp, _vp, _vp, _f32, _vp, _vp, _i64, _i32, _i32, _i32, _i32, _vp]),
    "b200lora_arena_adamw": (_i32, [_vp, _vp, _vp, _vp, _i64, _f32, _f32, _f32, _f32, _f32, _f32, _vp, _vp]),
    "b200lora_flat_axpby": (_i32, [_vp, _vp, _vp, _i64, _f32, _f32, _i32, _i32, _vp]),
    "b200lora_ti_grad_rows": (_i32, [_vp, _vp, _i64, _i32, _vp, _i32, _vp, _i32, _vp]),
    "b200lora_ti_step": (_i32, [_vp, _i32, _i32, _vp, _i32, _vp, _vp, _vp, _vp, _i32, _f32, _f32, _f32, _f32, _f32,
                                _f32, _f32, _vp, _vp, _vp, _vp]),
    "b200lora_linear_bwd_dx_fused": (_i32, [_vp, _vp, _i32, _vp, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _i32, _vp]),
    "b200lora_lora_grads_fused": (_i32, [_vp, _vp, _i32, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i32, _i32, _vp,
                                         _i32, _vp]),
    "b200lora_conv2d_repack_weight": (_i32, [_vp, _i32, _vp, _i32, _i32, _i32, _i32, _i32, _i32, _vp]),
    "b200lora_conv2d_pack_lora": (_i32, [_vp, _vp, _i32, _vp, _i32, _vp, _vp, _vp, _i32, _i32, _i32, _i32,
                                         _i32, _i32, _vp]),
    "b200lora_conv2d_gemm": (_i32, [_vp, _vp, _i32, _vp, _vp, _f32, _vp, _vp] + [_i32] * 15 + [_vp]),
    "b200lora_conv2d_zcat": (_i32, [_vp, _vp, _i32, _f32, _vp, _vp, _i64, _i32, _i32, _i32, _vp]),
    "b200lora_conv2d_grad_a": (_i32, [_vp, _vp, _f32, _vp] + [_i32] * 14 + [_vp]),
    "b200lora_merge": (_i32, [_vp, _vp, _vp, _vp, _f32, _i32, _i32, _i32, _i32, _i32, _vp]),
    "b200lora_merge_batch_workspace": (_i64, [_i32]),
    "b200lora_merge_batch_plan": (_i32, [_vp, _i32, _i32, _vp, _i64, _vp, _vp]),
    "b200lora_merge_batch_run": (_i32, [_vp, _i32, _i64, _i32, _f32, _i32, _i32, _vp]),
    "b200lora_absmean_accum": (_i32, [_vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp]),
    "b200lora_svd_epilogue": (_i32, [_vp, _vp, _vp, _f32, _i32, _i32, _i32, _i32, _vp]),
}


class MergeItem(ctypes.Structure):
    """b200lora_merge_item of include/b200lora.h."""
    _fields_ = [("Win", _vp), ("Wout", _vp), ("up", _vp), ("down", _vp),
                ("N", ctypes.c_int32), ("K", ctypes.c_int32), ("r", ctypes.c_int32), ("reserved", ctypes.c_int32)]


class B200LoraError(RuntimeError):
    pass


def lib():
    """Load (once) and return the shared library; raise loudly if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise B200LoraError(
                f"{LIB_PATH} is not built. Run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(or `python diffusion_finetuning_b200/build.py`). There is no fallback path.")
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def dtype_code(dt):
    try:
        return _DTYPE_CODE[dt]
    except KeyError:
        raise B200LoraError(f"unsupported dtype {dt}; b200lora handles fp32/fp16/bf16") from None


# number of successful C-ABI kernel launches since import (every entry point that computes is
# exactly one kernel launch); bench.py reports the delta over its timed region as `gpu_launches`
LAUNCHES = 0
_DEBUG_ERR = os.environ.get("B200LORA_DEBUG_ERR", "0") == "1"


def check(rc, what):
    global LAUNCHES
    if rc != 0:
        msg = lib().b200lora_last_error().decode(errors="replace")
        raise B200LoraError(f"{what} failed with code {rc}: {msg}")
    LAUNCHES += 1
    if _DEBUG_ERR:
        code = lib().b200lora_peek_cuda_error()
        if code != 0:
            raise B200LoraError(f"{what} returned 0 but left CUDA error {code} pending")


def stream_ptr(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def require_cuda(t, what):
    if not t.is_cuda:
        raise B200LoraError(
            f"{what}: tensor is on {t.device}; the b200lora hot path is CUDA-only and has no CPU "
            "fallback. Move the module to a B200 (`.to('cuda')`).")
