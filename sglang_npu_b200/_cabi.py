"""ctypes binding of the C-ABI in include/b200_hotpath.h.

This is the ONLY way the Python host layer reaches device code.  There is no CPU or
PyTorch fallback: if the shared library is missing `load()` raises, and every wrapper
raises on a non-zero status (RuntimeError, or NotImplementedError for B200_ERR_UNSUPPORTED —
mirroring TORCH_CHECK / TORCH_CHECK_NOT_IMPLEMENTED in the reference,
sgl-kernel/csrc/gemm/int8_gemm_kernel.cu:670-695, :742-743).
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_int64, c_size_t, c_void_p, POINTER, byref
from typing import Optional

import torch

_LIB: Optional[ctypes.CDLL] = None
LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib", "libb200_hotpath.so")

B200_OK = 0
B200_ERR_INVALID_ARG = 1
B200_ERR_UNSUPPORTED = 2
B200_ERR_CUDA = 3
B200_ERR_WORKSPACE = 4

DTYPE_CODE = {
    torch.bfloat16: 0,
    torch.float16: 1,
    torch.float32: 2,
    torch.float8_e4m3fn: 3,
    torch.int8: 4,
}

# name -> (argtypes); every entry point returns int.  Must list every symbol of the header
# (tests/test_cabi_symbols.py parses include/b200_hotpath.h and checks this table against it).
_P = c_void_p
_SIGNATURES = {
    "b200_abi_version": [],
    "b200_device_info": [POINTER(c_int), POINTER(c_int), POINTER(c_int)],
    "b200_debug_set_trace": [_P, c_size_t],
    "b200_set_kv_buffer": [_P, _P, _P, _P, _P, c_int, c_int64, c_int64, c_int64, c_int64, c_int64,
                           c_int64, c_int64, c_int, _P],
    "b200_set_kv_buffer_fp8": [_P, _P, _P, _P, _P, c_int, c_int64, c_int64, c_int64, c_int64, c_int64,
                               c_int64, c_int64, c_float, c_float, c_int, _P],
    "b200_fp8_kv_gather_dequant": [_P, _P, _P, _P, _P, c_int64, _P, c_int, _P, _P, c_int, c_int, c_int64, c_int64,
                                   c_float, c_float, c_int, _P],
    "b200_decode_attn_workspace_bytes": [c_int, c_int, c_int, c_int, POINTER(c_size_t)],
    "b200_decode_attn": [_P, _P, _P, _P, _P, c_int64, _P, _P, c_int, c_int, c_int, c_int, c_int,
                         c_int64, c_int64, c_int64, c_float, c_float, c_int, c_int, _P, c_size_t,
                         c_int, _P],
    "b200_decode_attn_int8": [_P, _P, _P, _P, _P, _P, _P, c_int64, _P, _P, c_int, c_int, c_int, c_int, c_int,
                              c_int64, c_int64, c_int64, c_float, c_float, c_int, c_int, _P, c_size_t,
                              c_int, _P],
    "b200_decode_attn_fp8kv": [_P, _P, _P, _P, _P, _P, _P, c_int64, _P, _P, c_int, c_int, c_int, c_int, c_int,
                               c_int64, c_int64, c_int64, c_float, c_float, c_float, c_float, c_int, c_int, _P,
                               c_size_t, c_int, _P],
    "b200_decode_attn_shared_prefix": [_P, _P, _P, _P, _P, _P, _P, c_int64, _P, _P, c_int, c_int, c_int, c_int, c_int,
                                       c_int64, c_int64, c_int64, c_float, c_float, c_int, c_int, c_int, c_int, _P,
                                       c_size_t, c_int, _P, c_int64],
    "b200_decode_attn_shared_prefix_fp8kv": [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, c_int64, _P, c_int64, _P, _P, c_int,
                                             c_int, c_int, c_int, c_int, c_int64, c_int64, c_int64, c_float, c_float,
                                             c_float, c_float, c_int, c_int, c_int, c_int, _P, c_size_t, c_int, _P],
    "b200_rmsnorm_group_quant_fp8": [_P, _P, _P, _P, _P, c_int64, c_int64, c_float, c_float, c_int, _P],
    "b200_silu_and_mul_group_quant_fp8": [_P, _P, _P, c_int64, c_int64, c_float, c_int, _P],
    "b200_extend_attn": [_P, _P, _P, _P, _P, _P, _P, c_int64, _P, c_int, _P, _P, c_int, c_int,
                         c_int, c_int, c_int, c_int64, c_int64, c_int64, c_int64, c_int64, c_float,
                         c_float, c_int, c_int, _P, c_int64, c_int64],
    "b200_merge_state": [_P, _P, _P, _P, _P, _P, c_int64, c_int, c_int, c_int, _P],
    "b200_per_token_quant_int8": [_P, _P, _P, c_int64, c_int64, c_int64, c_int, _P],
    "b200_int8_scaled_mm_workspace_bytes": [c_int64, c_int64, c_int64, POINTER(c_size_t)],
    "b200_int8_scaled_mm": [_P, _P, _P, _P, _P, _P, c_int64, c_int64, c_int64, c_int, _P, _P,
                            c_size_t, _P],
    "b200_per_token_group_quant_fp8": [_P, _P, _P, c_int64, c_int64, c_int64, c_int, c_float,
                                       c_int64, c_int64, c_int, _P],
    "b200_per_token_quant_fp8": [_P, _P, _P, c_int64, c_int64, c_int64, c_int, _P],
    "b200_fp8_blockwise_scaled_mm": [_P, _P, _P, _P, c_int64, c_int64, _P, c_int64, c_int64, _P,
                                     c_int64, c_int64, c_int64, c_int, _P, c_size_t, _P],
    "b200_fp8_scaled_mm": [_P, _P, _P, _P, _P, _P, c_int64, c_int64, c_int64, c_int, _P, c_size_t, _P],
    "b200_awq_dequantize": [_P, _P, _P, _P, c_int64, c_int64, c_int, c_int, _P],
    "b200_awq_repack": [_P, _P, _P, _P, _P, c_int64, c_int64, c_int, c_int, _P],
    "b200_awq_gemm_workspace_bytes": [POINTER(c_size_t)],
    "b200_awq_gemm": [_P, _P, _P, _P, _P, c_int64, c_int64, c_int64, c_int, c_int, _P, c_size_t, _P],
    "b200_rmsnorm": [_P, _P, _P, _P, c_int64, c_int64, c_float, c_int, _P],
    "b200_silu_and_mul": [_P, _P, c_int64, c_int64, c_int, _P],
    "b200_rope_neox": [_P, _P, _P, _P, c_int64, c_int, c_int, c_int, c_int, c_int64, c_int64,
                       c_int, _P],
    "b200_argmax_workspace_bytes": [c_int64, POINTER(c_size_t)],
    "b200_argmax": [_P, _P, c_int64, c_int64, c_int64, c_int, _P, c_size_t, _P],
    "b200_argmax_static": [_P, _P, c_int64, c_int64, c_int64, c_int, _P],
    "b200_argmax_key": [_P, _P, c_int64, c_int64, c_int64, c_int, c_int64, _P, c_size_t, _P],
    "b200_argmax_key_decode": [_P, _P, c_int64, _P],
    "b200_embedding": [_P, _P, _P, c_int64, c_int64, c_int64, c_int, _P, _P, _P],
    "b200_dense_gemm_nt": [_P, _P, _P, _P, c_int64, c_int64, c_int64, c_int, _P, c_size_t, _P],
    "b200_rmsnorm_quant_int8": [_P, _P, _P, _P, _P, _P, c_int64, c_int64, c_float, c_int, _P],
    "b200_silu_and_mul_quant_int8": [_P, _P, _P, c_int64, c_int64, c_int, _P],
    "b200_rope_neox_set_kv": [_P, _P, _P, _P, _P, _P, _P, _P, c_int, c_int64, c_int, c_int, c_int, c_int,
                              c_int64, c_int64, c_int64, c_int64, c_int64, c_int, _P],
    "b200_alloc_extend": [_P, _P, _P, _P, _P, c_int, _P, c_int, c_int, _P, c_int64, _P, c_int, c_int, _P],
    "b200_alloc_decode": [_P, _P, _P, _P, c_int, _P, c_int, c_int, _P, c_int64, _P, c_int, c_int, _P],
    "b200_write_req_to_token_pool": [_P, c_int64, _P, c_int, _P, _P, _P, c_int, _P, c_int, c_int, _P],
    "b200_get_last_loc": [_P, _P, c_int64, _P, c_int, _P, c_int, c_int, _P],
    "b200_compute_position": [_P, _P, _P, _P, c_int, c_int, _P],
    "b200_create_kv_indices": [_P, _P, c_int64, _P, c_int, _P, c_int, _P, _P, c_int, _P],
    "b200_comm_malloc": [POINTER(c_void_p), c_size_t],
    "b200_comm_free": [_P],
    "b200_comm_ipc_handle": [_P, _P],
    "b200_comm_ipc_open": [_P, POINTER(c_void_p)],
    "b200_comm_ipc_close": [_P],
    "b200_tp_reduce_rmsnorm_quant_int8": [_P, _P, _P, _P, _P, _P, _P, _P, c_int64, c_int64, c_float, c_int, c_int,
                                          c_int, c_int, _P, _P],
    "b200_tp_all_reduce": [_P, _P, _P, _P, c_int64, c_int64, c_int, c_int, c_int, c_int, _P, _P],
    "b200_tp_exchange_loopback": [_P, _P, _P, _P, _P, _P, _P, _P, c_int64, c_int64, c_float, c_int, c_int, c_int, _P, _P,
                                  c_size_t, _P],
    "b200_comm_check": [_P, _P],
}

EXPORTED_SYMBOLS = sorted(list(_SIGNATURES) + ["b200_last_error"])


def load() -> ctypes.CDLL:
    """Load libb200_hotpath.so (built in-tree by sglang_npu_b200.build).  Raises if absent."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: the B200 hot path has no CPU/PyTorch fallback. "
            "Build it with `python -m sglang_npu_b200.build` (or __graft_entry__.build())."
        )
    lib = ctypes.CDLL(LIB_PATH)
    for name, argtypes in _SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.argtypes = argtypes
        fn.restype = c_int
    lib.b200_last_error.argtypes = []
    lib.b200_last_error.restype = c_char_p
    _LIB = lib
    return lib


def check(status: int, what: str) -> None:
    if status == B200_OK:
        return
    msg = load().b200_last_error().decode("utf-8", "replace")
    if status == B200_ERR_UNSUPPORTED:
        raise NotImplementedError(f"{what}: {msg}")
    raise RuntimeError(f"{what} failed (status {status}): {msg}")


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def stream_ptr(device=None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def require_cuda(*tensors: torch.Tensor) -> None:
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError(
                "b200 hot path called with a non-CUDA tensor; there is no CPU fallback in this backend"
            )


def dtype_code(dt: torch.dtype) -> int:
    try:
        return DTYPE_CODE[dt]
    except KeyError:
        raise NotImplementedError(f"b200 hot path: dtype {dt} not supported") from None
