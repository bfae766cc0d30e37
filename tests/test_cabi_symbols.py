"""CPU: the C-ABI library builds, loads, and exports every symbol include/b200_hotpath.h declares
(no compute calls — there is no GPU here)."""
import os
import re
import subprocess

from sglang_npu_b200 import _cabi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200_hotpath.h")


def header_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_minimum_set():
    syms = header_symbols()
    for must in ["b200_set_kv_buffer", "b200_decode_attn", "b200_extend_attn", "b200_merge_state",
                 "b200_per_token_quant_int8", "b200_int8_scaled_mm", "b200_per_token_group_quant_fp8",
                 "b200_fp8_blockwise_scaled_mm", "b200_awq_gemm", "b200_awq_dequantize"]:
        assert must in syms


def test_library_exports_every_header_symbol(lib):
    out = subprocess.run(["nm", "-D", "--defined-only", _cabi.LIB_PATH], capture_output=True, text=True, check=True)
    exported = set(re.findall(r"\bT (b200_[a-z0-9_]+)", out.stdout))
    missing = [s for s in header_symbols() if s not in exported]
    assert not missing, f"declared in the header but not exported: {missing}"
    # and the ctypes table covers the header 1:1
    assert sorted(_cabi.EXPORTED_SYMBOLS) == header_symbols()


def test_abi_version_and_error_channel(lib):
    assert lib.b200_abi_version() == 2
    # argument validation happens before any CUDA call, so it is testable without a GPU
    st = lib.b200_set_kv_buffer(None, None, None, None, None, 0, 4, 8, 8, 8, 8, 8, 8, 2, None)
    assert st == _cabi.B200_ERR_INVALID_ARG
    assert b"null pointer" in lib.b200_last_error()


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    import pytest

    monkeypatch.setattr(_cabi, "_LIB", None)
    monkeypatch.setattr(_cabi, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(RuntimeError, match="no CPU/PyTorch fallback"):
        _cabi.load()


def test_cpu_tensors_are_rejected():
    import pytest
    import torch

    from sglang_npu_b200 import kernels

    with pytest.raises(RuntimeError, match="no CPU fallback"):
        kernels.per_token_quant_int8(torch.zeros(2, 16, dtype=torch.bfloat16))


def test_argmax_static_rejects_oversize_before_touching_cuda(lib):
    """b200_argmax_static (scratch in a static device array, 4096 rows): the size check comes before any CUDA call."""
    st = lib.b200_argmax_static(None, None, 4097, 128, 128, 1, None)
    assert st == _cabi.B200_ERR_WORKSPACE
    assert b"4096" in lib.b200_last_error()
    assert lib.b200_argmax_static(None, None, 0, 128, 128, 1, None) == 0
