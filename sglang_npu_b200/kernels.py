"""Python op wrappers over the C-ABI — the replacement for `sgl_kernel`'s Python layer
(reference sgl-kernel/python/sgl_kernel/{gemm,attention,elementwise}.py).

Names and argument meaning follow the reference ops so the layer code reads the same;
outputs are allocated here with torch (ownership contract of SURVEY §8b) and the kernels
are enqueued on the current CUDA stream.  No fallbacks: non-CUDA tensors raise.
"""
from __future__ import annotations

import os
from ctypes import byref, c_size_t
from typing import Optional, Tuple

import torch

from . import _cabi
from ._cabi import check as _check, dtype_code, ptr, require_cuda, stream_ptr


_LAUNCHES = 0
_KERNELS_PER_CALL = {"b200_decode_attn_shared_prefix": 3, "b200_decode_attn_shared_prefix_fp8kv": 3, "b200_decode_attn": 2, "b200_decode_attn_int8": 2, "b200_decode_attn_fp8kv": 2, "b200_decode_attn_workspace_bytes": 0,
                     "b200_int8_scaled_mm_workspace_bytes": 0, "b200_argmax_workspace_bytes": 0}


def check(status: int, what: str) -> None:
    _check(status, what)
    _count(_KERNELS_PER_CALL.get(what, 1))



def reset_launch_count() -> None:
    global _LAUNCHES
    _LAUNCHES = 0


def launch_count() -> int:
    """Number of kernels of THIS library launched through the wrappers since the last reset."""
    return _LAUNCHES


def _count(n: int = 1) -> None:
    global _LAUNCHES
    _LAUNCHES += n


def set_trace(buf: Optional[torch.Tensor]) -> None:
    """Diagnostics: per-CTA phase timestamps of the GEMM kernels into `buf` (uint64/int64 [ctas, 16]); None = off."""
    lib = _cabi.load()
    _check(lib.b200_debug_set_trace(ptr(buf), 0 if buf is None else buf.numel() * buf.element_size()),
           "b200_debug_set_trace")


def _is64(t: torch.Tensor) -> int:
    if t.dtype == torch.int64:
        return 1
    if t.dtype == torch.int32:
        return 0
    raise RuntimeError(f"index tensor must be int32 or int64, got {t.dtype}")


# ---- KV write ------------------------------------------------------------------------------
_FP8_STORE = (torch.float8_e4m3fn, torch.uint8)


def set_kv_buffer(k_cache: torch.Tensor, v_cache: torch.Tensor, loc: torch.Tensor,
                  cache_k: torch.Tensor, cache_v: torch.Tensor, k_scale: Optional[float] = None,
                  v_scale: Optional[float] = None) -> None:
    """k_cache[loc] = cache_k; v_cache[loc] = cache_v (memory_pool.py:369-407), one launch.  When the cache is FP8
    (float8_e4m3fn, or its uint8 store view) and cache_k/v are bf16/f16, the cast branch (:385-391: optional division
    by k_scale / v_scale, then `.to(float8_e4m3fn)`) runs inside the same launch."""
    require_cuda(k_cache, v_cache, loc, cache_k, cache_v)
    lib = _cabi.load()
    T = loc.numel()
    if T == 0:
        return
    assert cache_k.shape[0] == T and cache_v.shape[0] == T
    if cache_k.dtype != k_cache.dtype and k_cache.dtype in _FP8_STORE and cache_k.dtype in (torch.bfloat16, torch.float16):
        assert cache_v.dtype == cache_k.dtype and v_cache.dtype in _FP8_STORE
        cache_k = cache_k.reshape(T, -1)
        cache_v = cache_v.reshape(T, -1)
        cache_k = cache_k if cache_k.stride(-1) == 1 else cache_k.contiguous()
        cache_v = cache_v if cache_v.stride(-1) == 1 else cache_v.contiguous()
        loc = loc if loc.is_contiguous() else loc.contiguous()
        kc = k_cache.view(k_cache.shape[0], -1)
        vc = v_cache.view(v_cache.shape[0], -1)
        assert kc.shape[1] == cache_k.shape[1] and vc.shape[1] == cache_v.shape[1]
        check(lib.b200_set_kv_buffer_fp8(
            ptr(kc), ptr(vc), ptr(cache_k), ptr(cache_v), ptr(loc), _is64(loc), T,
            cache_k.shape[1], cache_v.shape[1], cache_k.stride(0), cache_v.stride(0), kc.stride(0), vc.stride(0),
            float(k_scale) if k_scale is not None else 0.0, float(v_scale) if v_scale is not None else 0.0,
            dtype_code(cache_k.dtype), stream_ptr(k_cache.device)), "b200_set_kv_buffer_fp8")
        return
    assert cache_k.dtype == k_cache.dtype and cache_v.dtype == v_cache.dtype
    cache_k = cache_k.reshape(T, -1) if cache_k.dim() != 2 else cache_k
    cache_v = cache_v.reshape(T, -1) if cache_v.dim() != 2 else cache_v
    if cache_k.stride(-1) != 1:
        cache_k = cache_k.contiguous()
    if cache_v.stride(-1) != 1:
        cache_v = cache_v.contiguous()
    if not loc.is_contiguous():
        loc = loc.contiguous()
    kc = k_cache.view(k_cache.shape[0], -1)
    vc = v_cache.view(v_cache.shape[0], -1)
    assert kc.shape[1] == cache_k.shape[1] and vc.shape[1] == cache_v.shape[1]
    check(lib.b200_set_kv_buffer(
        ptr(kc), ptr(vc), ptr(cache_k), ptr(cache_v), ptr(loc), _is64(loc), T,
        cache_k.shape[1], cache_v.shape[1], cache_k.stride(0), cache_v.stride(0),
        kc.stride(0), vc.stride(0), k_cache.element_size(), stream_ptr(k_cache.device)),
        "b200_set_kv_buffer")


def fp8_kv_gather_dequant(k_out: torch.Tensor, v_out: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                          req_to_token: torch.Tensor, req_pool_indices: torch.Tensor, prefix_lens: torch.Tensor,
                          offsets: torch.Tensor, max_prefix_len: int, k_scale: Optional[float] = None,
                          v_scale: Optional[float] = None) -> None:
    """k_out[offsets[b] + pos] = up-cast of the fp8 cache row of prefix token `pos` of request b (and v likewise);
    k_out/v_out [rows, Hkv, D] bf16/f16.  Staging for extend attention over an FP8 KV cache."""
    require_cuda(k_out, v_out, k_cache, v_cache, req_to_token, req_pool_indices, prefix_lens, offsets)
    lib = _cabi.load()
    assert k_cache.dtype in _FP8_STORE and v_cache.dtype in _FP8_STORE and k_out.dtype == v_out.dtype
    assert prefix_lens.dtype == torch.int32 and offsets.dtype == torch.int32 and req_to_token.dtype == torch.int32
    assert k_out.is_contiguous() and v_out.is_contiguous()
    kc = k_cache.view(k_cache.shape[0], -1)
    vc = v_cache.view(v_cache.shape[0], -1)
    assert kc.stride(1) == 1 and kc.stride() == vc.stride() and kc.shape[1] == k_out[0].numel()
    check(lib.b200_fp8_kv_gather_dequant(
        ptr(k_out), ptr(v_out), ptr(kc), ptr(vc), ptr(req_to_token), req_to_token.stride(0), ptr(req_pool_indices),
        _is64(req_pool_indices), ptr(prefix_lens), ptr(offsets), req_pool_indices.numel(), int(max_prefix_len),
        kc.shape[1], kc.stride(0), 1.0 if k_scale is None else float(k_scale),
        1.0 if v_scale is None else float(v_scale), dtype_code(k_out.dtype), stream_ptr(k_out.device)),
        "b200_fp8_kv_gather_dequant")


# ---- decode attention -------------------------------------------------------------------------
def decode_attn_workspace_bytes(batch: int, num_q_heads: int, head_dim: int, max_splits: int) -> int:
    lib = _cabi.load()
    out = c_size_t(0)
    check(lib.b200_decode_attn_workspace_bytes(batch, num_q_heads, head_dim, max_splits, byref(out)),
          "b200_decode_attn_workspace_bytes")
    return out.value


def decode_attention_fwd(q: torch.Tensor, k_buffer: torch.Tensor, v_buffer: torch.Tensor,
                         o: torch.Tensor, req_to_token: torch.Tensor,
                         req_pool_indices: torch.Tensor, seq_lens: torch.Tensor,
                         workspace: Optional[torch.Tensor], kv_chunk: int, max_splits: int,
                         sm_scale: float, logit_cap: float = 0.0, int8_out: bool = False,
                         k_scale: Optional[float] = None, v_scale: Optional[float] = None):
    """q [bs, Hq, D]; k/v_buffer [slots, Hkv, D]; o [bs, Hq, D] (written).  int8_out=True fuses the per-token int8
    quantisation of the [bs, Hq*D] output into stage 2 and returns it as an Int8Activation (for o_proj).  An FP8
    (float8_e4m3fn / uint8 store view) k/v_buffer selects the FP8-KV kernel; k_scale / v_scale apply only there."""
    require_cuda(q, k_buffer, v_buffer, o, req_to_token, req_pool_indices, seq_lens)
    lib = _cabi.load()
    bs, hq, d = q.shape
    hkv = k_buffer.shape[1]
    assert q.stride(2) == 1 and q.stride(1) == d and o.stride(2) == 1 and o.stride(1) == d
    assert k_buffer.stride(2) == 1 and k_buffer.stride(1) == d and v_buffer.stride() == k_buffer.stride()
    assert req_to_token.dtype == torch.int32 and req_to_token.stride(1) == 1
    assert req_pool_indices.dtype == seq_lens.dtype
    assert req_pool_indices.is_contiguous() and seq_lens.is_contiguous()
    ws_bytes = 0 if workspace is None else workspace.numel() * workspace.element_size()
    if k_buffer.dtype in _FP8_STORE:
        assert v_buffer.dtype in _FP8_STORE
        oq = torch.empty((bs, hq * d), device=q.device, dtype=torch.int8) if int8_out else None
        oscale = torch.empty((bs, 1), device=q.device, dtype=torch.float32) if int8_out else None
        check(lib.b200_decode_attn_fp8kv(
            ptr(o), ptr(oq), ptr(oscale), ptr(q), ptr(k_buffer), ptr(v_buffer), ptr(req_to_token),
            req_to_token.stride(0), ptr(req_pool_indices), ptr(seq_lens), _is64(seq_lens), bs, hq, hkv, d,
            q.stride(0), o.stride(0), k_buffer.stride(0), float(sm_scale), float(logit_cap),
            1.0 if k_scale is None else float(k_scale), 1.0 if v_scale is None else float(v_scale),
            int(kv_chunk), int(max_splits), ptr(workspace), ws_bytes, dtype_code(q.dtype),
            stream_ptr(q.device)), "b200_decode_attn_fp8kv")
        return Int8Activation(oq, oscale, q.dtype) if int8_out else None
    assert k_buffer.dtype == q.dtype
    if int8_out:
        oq = torch.empty((bs, hq * d), device=q.device, dtype=torch.int8)
        oscale = torch.empty((bs, 1), device=q.device, dtype=torch.float32)
        check(lib.b200_decode_attn_int8(
            ptr(o), ptr(oq), ptr(oscale), ptr(q), ptr(k_buffer), ptr(v_buffer), ptr(req_to_token),
            req_to_token.stride(0), ptr(req_pool_indices), ptr(seq_lens), _is64(seq_lens), bs, hq, hkv, d,
            q.stride(0), o.stride(0), k_buffer.stride(0), float(sm_scale), float(logit_cap),
            int(kv_chunk), int(max_splits), ptr(workspace), ws_bytes, dtype_code(q.dtype),
            stream_ptr(q.device)), "b200_decode_attn_int8")
        return Int8Activation(oq, oscale, q.dtype)
    check(lib.b200_decode_attn(
        ptr(o), ptr(q), ptr(k_buffer), ptr(v_buffer), ptr(req_to_token), req_to_token.stride(0),
        ptr(req_pool_indices), ptr(seq_lens), _is64(seq_lens), bs, hq, hkv, d,
        q.stride(0), o.stride(0), k_buffer.stride(0), float(sm_scale), float(logit_cap),
        int(kv_chunk), int(max_splits), ptr(workspace), ws_bytes, dtype_code(q.dtype),
        stream_ptr(q.device)), "b200_decode_attn")


def plan_shared_prefix_segments(shared_prefix_len: int, batch: int, num_q_heads: int, sm_count: int = 148,
                                sm_fraction: Optional[float] = None, head_dim: int = 128) -> int:
    """Segment length (multiple of 128) for the prefix pass of the shared-prefix decode.  The pass runs one CTA per SM
    (512 TMEM columns, 198 KB smem), ceil(batch/128) * Hq CTAs per segment: take as many segments as fit in ONE wave
    (2.05 waves measured 0.140 ms vs 0.122 ms for one), at least 512 tokens each.  Leaving SMs free for the forked
    per-request pass does not pay (cfg 5 per GPU: sm_fraction 0.5 / 0.65 / 0.8 / 1.0 -> 0.160 / 0.131 / 0.125 /
    0.117 ms): that pass needs every SM to saturate HBM, so the two passes' SM-time simply adds up."""
    if sm_fraction is None:
        sm_fraction = float(os.environ.get("B200_PREFIX_SM_FRACTION", "1.0"))
    # query rows per CTA of the prefix pass: 256 for the two-query-tile kernel (head_dim 128, more than one 128-row tile
    # of requests; extend_attn_fa.cu), else 128 (extend_attn_umma.cu) — launch_prefix_pass in csrc/decode_attn.cu
    rows = 256 if (batch > 128 and head_dim == 128 and os.environ.get("B200_EXTEND_FA", "1") != "0") else 128
    ctas_per_seg = max(1, (batch + rows - 1) // rows) * num_q_heads
    n_seg = max(1, int(sm_count * sm_fraction) // ctas_per_seg)
    seg = (shared_prefix_len + n_seg - 1) // n_seg
    return max(512, (seg + 127) // 128 * 128)


def decode_attention_shared_prefix_fwd(q: torch.Tensor, k_buffer: torch.Tensor, v_buffer: torch.Tensor,
                                       o: torch.Tensor, req_to_token: torch.Tensor, req_pool_indices: torch.Tensor,
                                       seq_lens: torch.Tensor, shared_prefix_len: int, seg_len: int,
                                       workspace: torch.Tensor, kv_chunk: int, max_splits: int, sm_scale: float,
                                       logit_cap: float = 0.0, int8_out: bool = False, stage=None,
                                       k_scale: Optional[float] = None, v_scale: Optional[float] = None):
    """Decode attention for a batch whose requests all start with the same `shared_prefix_len` KV slots (request 0's
    page-table row names them); same result as decode_attention_fwd.  See b200_decode_attn_shared_prefix.
    FP8 (e4m3) k/v_buffer: `stage` = (k_stage, v_stage, stage_table, stage_index) — the 16-bit staging copy of the
    shared tokens made by fp8_kv_gather_dequant (b200_decode_attn_shared_prefix_fp8kv)."""
    require_cuda(q, k_buffer, v_buffer, o, req_to_token, req_pool_indices, seq_lens, workspace)
    lib = _cabi.load()
    bs, hq, d = q.shape
    hkv = k_buffer.shape[1]
    assert q.stride(2) == 1 and q.stride(1) == d and o.stride(2) == 1 and o.stride(1) == d
    assert k_buffer.stride(2) == 1 and k_buffer.stride(1) == d and v_buffer.stride() == k_buffer.stride()
    assert req_to_token.dtype == torch.int32 and req_to_token.stride(1) == 1
    assert req_pool_indices.dtype == seq_lens.dtype
    assert req_pool_indices.is_contiguous() and seq_lens.is_contiguous()
    oq = torch.empty((bs, hq * d), device=q.device, dtype=torch.int8) if int8_out else None
    oscale = torch.empty((bs, 1), device=q.device, dtype=torch.float32) if int8_out else None
    if k_buffer.dtype in _FP8_STORE:
        assert v_buffer.dtype in _FP8_STORE and stage is not None, "FP8 KV: the staged shared prefix is required"
        k_stage, v_stage, stage_table, stage_index = stage
        require_cuda(k_stage, v_stage, stage_table, stage_index)
        assert k_stage.dtype == q.dtype and v_stage.dtype == q.dtype and k_stage.shape[0] >= shared_prefix_len
        assert k_stage.shape[1:] == (hkv, d) and k_stage.is_contiguous() and v_stage.is_contiguous()
        assert stage_table.dtype == torch.int32 and stage_table.numel() >= shared_prefix_len
        assert stage_index.dtype == req_pool_indices.dtype
        check(lib.b200_decode_attn_shared_prefix_fp8kv(
            ptr(o), ptr(oq), ptr(oscale), ptr(q), ptr(k_buffer), ptr(v_buffer), ptr(k_stage), ptr(v_stage),
            ptr(stage_table), ptr(stage_index), k_stage.stride(0), ptr(req_to_token), req_to_token.stride(0),
            ptr(req_pool_indices), ptr(seq_lens), _is64(seq_lens), bs, hq, hkv, d, q.stride(0), o.stride(0),
            k_buffer.stride(0), float(sm_scale), float(logit_cap), 1.0 if k_scale is None else float(k_scale),
            1.0 if v_scale is None else float(v_scale), int(shared_prefix_len), int(seg_len), int(kv_chunk),
            int(max_splits), ptr(workspace), workspace.numel() * workspace.element_size(), dtype_code(q.dtype),
            stream_ptr(q.device)), "b200_decode_attn_shared_prefix_fp8kv")
        return Int8Activation(oq, oscale, q.dtype) if int8_out else None
    assert k_buffer.dtype == q.dtype
    check(lib.b200_decode_attn_shared_prefix(
        ptr(o), ptr(oq), ptr(oscale), ptr(q), ptr(k_buffer), ptr(v_buffer), ptr(req_to_token),
        req_to_token.stride(0), ptr(req_pool_indices), ptr(seq_lens), _is64(seq_lens), bs, hq, hkv, d,
        q.stride(0), o.stride(0), k_buffer.stride(0), float(sm_scale), float(logit_cap), int(shared_prefix_len),
        int(seg_len), int(kv_chunk), int(max_splits), ptr(workspace),
        workspace.numel() * workspace.element_size(), dtype_code(q.dtype), stream_ptr(q.device), k_buffer.shape[0]),
        "b200_decode_attn_shared_prefix")
    return Int8Activation(oq, oscale, q.dtype) if int8_out else None


# ---- extend attention -------------------------------------------------------------------------
def extend_attention_fwd(q: torch.Tensor, k_new: Optional[torch.Tensor], v_new: Optional[torch.Tensor],
                         o: torch.Tensor, k_buffer: torch.Tensor, v_buffer: torch.Tensor,
                         req_to_token: torch.Tensor, req_pool_indices: torch.Tensor,
                         qo_indptr: torch.Tensor, prefix_lens: torch.Tensor, max_extend_len: int,
                         sm_scale: float, logit_cap: float = 0.0, causal: bool = True) -> None:
    """q [T, Hq, D]; k_new/v_new [T, Hkv, D] or None (read new keys from the cache)."""
    require_cuda(q, o, k_buffer, v_buffer, req_to_token, req_pool_indices, qo_indptr, prefix_lens)
    lib = _cabi.load()
    T, hq, d = q.shape
    hkv = k_buffer.shape[1]
    bs = req_pool_indices.numel()
    assert qo_indptr.dtype == torch.int32 and prefix_lens.dtype == torch.int32
    assert q.stride(2) == 1 and q.stride(1) == d and o.stride(2) == 1 and o.stride(1) == d
    kn_stride = vn_stride = 0
    if k_new is not None:
        assert k_new.stride(2) == 1 and k_new.stride(1) == d and v_new.stride(2) == 1 and v_new.stride(1) == d
        kn_stride, vn_stride = k_new.stride(0), v_new.stride(0)
    check(lib.b200_extend_attn(
        ptr(o), ptr(q), ptr(k_new), ptr(v_new), ptr(k_buffer), ptr(v_buffer), ptr(req_to_token),
        req_to_token.stride(0), ptr(req_pool_indices), _is64(req_pool_indices), ptr(qo_indptr),
        ptr(prefix_lens), bs, int(max_extend_len), hq, hkv, d, q.stride(0), o.stride(0),
        kn_stride, vn_stride, k_buffer.stride(0), float(sm_scale), float(logit_cap), int(causal),
        dtype_code(q.dtype), stream_ptr(q.device), T, k_buffer.shape[0]), "b200_extend_attn")


# ---- LSE merge ---------------------------------------------------------------------------------
def merge_state_v2(v_a: torch.Tensor, s_a: torch.Tensor, v_b: torch.Tensor, s_b: torch.Tensor,
                   v_merged: Optional[torch.Tensor] = None, s_merged: Optional[torch.Tensor] = None
                   ) -> Tuple[torch.Tensor, torch.Tensor]:
    """sgl_kernel.merge_state_v2 (sgl-kernel/python/sgl_kernel/attention.py)."""
    require_cuda(v_a, s_a, v_b, s_b)
    lib = _cabi.load()
    s_a = s_a.to(torch.float32).contiguous()
    s_b = s_b.to(torch.float32).contiguous()
    v_a, v_b = v_a.contiguous(), v_b.contiguous()
    if v_merged is None:
        v_merged = torch.empty_like(v_a)
    if s_merged is None:
        s_merged = torch.empty_like(s_a)
    T, H, D = v_a.shape
    check(lib.b200_merge_state(ptr(v_merged), ptr(s_merged), ptr(v_a), ptr(s_a), ptr(v_b), ptr(s_b),
                               T, H, D, dtype_code(v_a.dtype), stream_ptr(v_a.device)), "b200_merge_state")
    return v_merged, s_merged


merge_state = merge_state_v2


# ---- int8 ------------------------------------------------------------------------------------------
def per_token_quant_int8(x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """int8_kernel.py:53-83 — returns (x_q int8 like x, scales fp32 [..., 1])."""
    require_cuda(x)
    lib = _cabi.load()
    K = x.shape[-1]
    x2 = x.reshape(-1, K)
    if x2.stride(-1) != 1:
        x2 = x2.contiguous()
    M = x2.shape[0]
    x_q = torch.empty((M, K), device=x.device, dtype=torch.int8)
    scales = torch.empty((M, 1), device=x.device, dtype=torch.float32)
    check(lib.b200_per_token_quant_int8(ptr(x_q), ptr(scales), ptr(x2), M, K, x2.stride(0),
                                        dtype_code(x.dtype), stream_ptr(x.device)),
          "b200_per_token_quant_int8")
    return x_q.view(x.shape), scales.view(x.shape[:-1] + (1,))


_INT8_WS = {}


def _int8_workspace(device, nbytes: int) -> torch.Tensor:
    # One self-cleaning stream-K workspace (flags + per-CTA partial accumulators) per device: the scaled GEMMs of
    # one process are stream-ordered on the forward stream (as in the reference's single-threaded TP worker).
    key = device.index
    ws = _INT8_WS.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.zeros(max(nbytes, 1), device=device, dtype=torch.uint8)
        _INT8_WS[key] = ws
    return ws


def _scaled_mm_workspace(lib, device, M: int, N: int, K: int) -> torch.Tensor:
    need = c_size_t(0)
    check(lib.b200_int8_scaled_mm_workspace_bytes(M, N, K, byref(need)), "b200_int8_scaled_mm_workspace_bytes")
    # sized for the largest decode tile up front so a CUDA-graph capture never sees a re-allocation
    return _int8_workspace(device, max(need.value, _WS_MIN_BYTES))


_WS_MIN_BYTES = 1024 + 160 * 256 * 128 * 4


def int8_scaled_mm(mat_a: torch.Tensor, mat_b: torch.Tensor, scales_a: torch.Tensor,
                   scales_b: torch.Tensor, out_dtype: torch.dtype,
                   bias: Optional[torch.Tensor] = None, return_acc: bool = False, cluster_splitk: bool = False,
                   out: Optional[torch.Tensor] = None):
    """sgl_kernel.int8_scaled_mm (sgl-kernel/python/sgl_kernel/gemm.py:20-29).

    mat_a [M,K] int8 row-major; mat_b [K,N] int8 COLUMN-major (weight.t()); scales fp32.
    `cluster_splitk=True` passes no workspace, which selects the cluster split-K decode kernel instead of the
    persistent stream-K one (A/B and parity of both CUDA kernels; results are bit-identical).
    """
    require_cuda(mat_a, mat_b, scales_a, scales_b)
    lib = _cabi.load()
    assert mat_a.dtype == torch.int8 and mat_b.dtype == torch.int8
    if mat_a.dim() != 2 or mat_b.dim() != 2:
        raise RuntimeError("int8_scaled_mm: mat_a and mat_b must be 2-D")
    M, K = mat_a.shape
    if mat_b.shape[0] != K:
        raise RuntimeError("int8_scaled_mm: inner dimensions differ")
    N = mat_b.shape[1]
    if mat_a.stride(1) != 1 or mat_a.stride(0) != K:
        raise RuntimeError("int8_scaled_mm: mat_a must be row-major contiguous")
    if mat_b.stride(0) != 1 or mat_b.stride(1) != K:
        raise RuntimeError("int8_scaled_mm: mat_b must be column-major (pass weight.t())")
    if scales_a.numel() != M or scales_b.numel() != N:
        raise RuntimeError("int8_scaled_mm: scale sizes do not match M / N")
    scales_a = scales_a.to(torch.float32).contiguous()
    scales_b = scales_b.to(torch.float32).contiguous()
    if bias is not None:
        assert bias.numel() == N and bias.dtype == out_dtype
        bias = bias.contiguous()
    if out is None:
        out = torch.empty((M, N), device=mat_a.device, dtype=out_dtype)
    elif out.shape != (M, N) or out.dtype != out_dtype or not out.is_contiguous():
        raise RuntimeError("int8_scaled_mm: `out` must be a contiguous [M, N] tensor of out_dtype")
    acc = torch.empty((M, N), device=mat_a.device, dtype=torch.int32) if return_acc else None
    ws = None if cluster_splitk else _scaled_mm_workspace(lib, mat_a.device, M, N, K)
    check(lib.b200_int8_scaled_mm(ptr(out), ptr(mat_a), ptr(mat_b), ptr(scales_a), ptr(scales_b),
                                  ptr(bias), M, N, K, dtype_code(out_dtype), ptr(acc), ptr(ws),
                                  0 if ws is None else ws.numel(), stream_ptr(mat_a.device)), "b200_int8_scaled_mm")
    return (out, acc) if return_acc else out


def _out_buffer(out: Optional[torch.Tensor], M: int, N: int, dtype: torch.dtype, device, what: str) -> torch.Tensor:
    """`out=` of the GEMM wrappers: the symmetric buffer of the fused TP exchange for row-parallel layers."""
    if out is None:
        return torch.empty((M, N), device=device, dtype=dtype)
    if tuple(out.shape) != (M, N) or out.dtype != dtype or not out.is_contiguous():
        raise RuntimeError(f"{what}: `out` must be a contiguous [M, N] tensor of the output dtype")
    return out


def dense_gemm_nt(a: torch.Tensor, b: torch.Tensor, bias: Optional[torch.Tensor] = None,
                  out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """a [M, K] @ b [N, K]^T (+ bias) for bf16 / f16, decode-size M (<= 256): the lm_head GEMM
    (logits_processor.py:470-500) on the weight-streaming tcgen05 kernel instead of a library call."""
    require_cuda(a, b, bias)
    lib = _cabi.load()
    if a.dim() != 2 or b.dim() != 2 or a.shape[1] != b.shape[1] or a.dtype != b.dtype:
        raise RuntimeError("dense_gemm_nt: a [M, K] and b [N, K] of one dtype expected")
    if a.stride(1) != 1 or a.stride(0) != a.shape[1] or b.stride(1) != 1 or b.stride(0) != b.shape[1]:
        raise RuntimeError("dense_gemm_nt: operands must be row-major contiguous")
    M, K = a.shape
    N = b.shape[0]
    out = _out_buffer(out, M, N, a.dtype, a.device, "dense_gemm_nt")
    if bias is not None:
        bias = bias.to(a.dtype).contiguous()
    ws = _scaled_mm_workspace(lib, a.device, M, N, K)
    check(lib.b200_dense_gemm_nt(ptr(out), ptr(a), ptr(b), ptr(bias), M, N, K, dtype_code(a.dtype), ptr(ws), ws.numel(),
                                 stream_ptr(a.device)), "b200_dense_gemm_nt")
    return out


# ---- fp8 ---------------------------------------------------------------------------------------------
def sgl_per_token_group_quant_fp8(x: torch.Tensor, group_size: int = 128, eps: float = 1e-10,
                                  column_major_scales: bool = False) -> Tuple[torch.Tensor, torch.Tensor]:
    """per_token_group_quant_fp8 (fp8_kernel.py:209-297): returns (x_q e4m3, x_s fp32 [M, K/G])."""
    require_cuda(x)
    lib = _cabi.load()
    K = x.shape[-1]
    x2 = x.reshape(-1, K)
    if x2.stride(-1) != 1:
        x2 = x2.contiguous()
    M = x2.shape[0]
    G = K // group_size
    x_q = torch.empty((M, K), device=x.device, dtype=torch.float8_e4m3fn)
    if column_major_scales:
        x_s = torch.empty((G, M), device=x.device, dtype=torch.float32).t()
    else:
        x_s = torch.empty((M, G), device=x.device, dtype=torch.float32)
    check(lib.b200_per_token_group_quant_fp8(ptr(x_q), ptr(x_s), ptr(x2), M, K, x2.stride(0), group_size,
                                             float(eps), x_s.stride(0), x_s.stride(1), dtype_code(x.dtype),
                                             stream_ptr(x.device)), "b200_per_token_group_quant_fp8")
    return x_q.view(x.shape), x_s.view(x.shape[:-1] + (G,)) if not column_major_scales else x_s


def sgl_per_token_quant_fp8(x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    require_cuda(x)
    lib = _cabi.load()
    K = x.shape[-1]
    x2 = x.reshape(-1, K)
    if x2.stride(-1) != 1:
        x2 = x2.contiguous()
    M = x2.shape[0]
    x_q = torch.empty((M, K), device=x.device, dtype=torch.float8_e4m3fn)
    x_s = torch.empty((M, 1), device=x.device, dtype=torch.float32)
    check(lib.b200_per_token_quant_fp8(ptr(x_q), ptr(x_s), ptr(x2), M, K, x2.stride(0), dtype_code(x.dtype),
                                       stream_ptr(x.device)), "b200_per_token_quant_fp8")
    return x_q.view(x.shape), x_s


def fp8_blockwise_scaled_mm(mat_a: torch.Tensor, mat_b: torch.Tensor, scales_a: torch.Tensor,
                            scales_b: torch.Tensor, out_dtype: torch.dtype,
                            bias: Optional[torch.Tensor] = None, cluster_splitk: bool = False,
                            out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """sgl_kernel.fp8_blockwise_scaled_mm (gemm.py:32-42): mat_a [M,K] e4m3 row-major, mat_b [K,N]
    e4m3 column-major (weight.T), scales_a [M, K/128], scales_b [K/128, N/128] (weight_scale.T)."""
    require_cuda(mat_a, mat_b, scales_a, scales_b)
    lib = _cabi.load()
    M, K = mat_a.shape
    N = mat_b.shape[1]
    if mat_b.shape[0] != K or mat_b.stride(0) != 1 or mat_b.stride(1) != K:
        raise RuntimeError("fp8_blockwise_scaled_mm: mat_b must be [K,N] column-major")
    if mat_a.stride(1) != 1 or mat_a.stride(0) != K:
        raise RuntimeError("fp8_blockwise_scaled_mm: mat_a must be row-major contiguous")
    if scales_a.shape != (M, K // 128) or scales_b.shape != (K // 128, N // 128):
        raise RuntimeError("fp8_blockwise_scaled_mm: scale shapes must be [M,K/128] and [K/128,N/128]")
    assert scales_a.dtype == torch.float32 and scales_b.dtype == torch.float32
    out = _out_buffer(out, M, N, out_dtype, mat_a.device, "fp8_blockwise_scaled_mm")
    if bias is not None:
        bias = bias.to(out_dtype).contiguous()
    # `cluster_splitk=True` passes no workspace: decode sizes then take the cluster split-K kernel instead of the
    # stream-K one (A/B and parity of both CUDA kernels)
    ws = None if cluster_splitk else _scaled_mm_workspace(lib, mat_a.device, M, N, K)
    # kernel wants scale_b indexed [n_block, k_block]
    check(lib.b200_fp8_blockwise_scaled_mm(
        ptr(out), ptr(mat_a), ptr(mat_b), ptr(scales_a), scales_a.stride(0), scales_a.stride(1),
        ptr(scales_b), scales_b.stride(1), scales_b.stride(0), ptr(bias), M, N, K,
        dtype_code(out_dtype), ptr(ws), 0 if ws is None else ws.numel(), stream_ptr(mat_a.device)),
        "b200_fp8_blockwise_scaled_mm")
    return out


def fp8_scaled_mm(mat_a: torch.Tensor, mat_b: torch.Tensor, scales_a: torch.Tensor,
                  scales_b: torch.Tensor, out_dtype: torch.dtype,
                  bias: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """sgl_kernel.fp8_scaled_mm: per-token x per-channel scaled e4m3 GEMM."""
    require_cuda(mat_a, mat_b, scales_a, scales_b)
    lib = _cabi.load()
    M, K = mat_a.shape
    N = mat_b.shape[1]
    if mat_b.shape[0] != K or mat_b.stride(0) != 1 or mat_b.stride(1) != K:
        raise RuntimeError("fp8_scaled_mm: mat_b must be [K,N] column-major")
    scales_a = scales_a.to(torch.float32).contiguous()
    scales_b = scales_b.to(torch.float32).contiguous()
    out = _out_buffer(out, M, N, out_dtype, mat_a.device, "fp8_scaled_mm")
    if bias is not None:
        bias = bias.to(out_dtype).contiguous()
    ws = _scaled_mm_workspace(lib, mat_a.device, M, N, K)
    check(lib.b200_fp8_scaled_mm(ptr(out), ptr(mat_a), ptr(mat_b), ptr(scales_a), ptr(scales_b), ptr(bias),
                                 M, N, K, dtype_code(out_dtype), ptr(ws), ws.numel(), stream_ptr(mat_a.device)),
          "b200_fp8_scaled_mm")
    return out


# ---- AWQ -----------------------------------------------------------------------------------------------
def awq_dequantize(qweight: torch.Tensor, scales: torch.Tensor, qzeros: torch.Tensor) -> torch.Tensor:
    """sgl_kernel.awq_dequantize (gemm.py:7-10): returns W [K, N] in the dtype of `scales`."""
    require_cuda(qweight, scales, qzeros)
    lib = _cabi.load()
    K, n8 = qweight.shape
    N = n8 * 8
    G = K // qzeros.shape[0]
    assert qweight.dtype == torch.int32 and qzeros.dtype == torch.int32
    assert qweight.is_contiguous() and qzeros.is_contiguous() and scales.is_contiguous()
    out = torch.empty((K, N), device=qweight.device, dtype=scales.dtype)
    check(lib.b200_awq_dequantize(ptr(out), ptr(qweight), ptr(scales), ptr(qzeros), K, N, G,
                                  dtype_code(scales.dtype), stream_ptr(qweight.device)), "b200_awq_dequantize")
    return out


_AWQ_WS = {}


def _awq_workspace(lib, device) -> torch.Tensor:
    """Stream-K state of the fused AWQ GEMM (launch epoch, publish flags, partial accumulators), one zero-initialised
    buffer per device; the AWQ GEMMs of a process are stream-ordered on the forward stream."""
    ws = _AWQ_WS.get(device.index)
    if ws is None:
        need = c_size_t(0)
        check(lib.b200_awq_gemm_workspace_bytes(byref(need)), "b200_awq_gemm_workspace_bytes")
        ws = torch.zeros(need.value, device=device, dtype=torch.uint8)
        _AWQ_WS[device.index] = ws
    return ws


def awq_repack(qweight: torch.Tensor, scales: torch.Tensor, qzeros: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Checkpoint layout (qweight [K, N/8], scales [K/g, N], qzeros [K/g, N/8]) -> the streaming layout of the fused GEMM:
    (qweight_t [units, 256, 16] int32 — one 16 KB image per (256-channel tile, 128-k block) — and sz [K/g, N] int32 =
    {scale, magic + zero}).  One-time, at process_weights_after_loading."""
    require_cuda(qweight, scales, qzeros)
    lib = _cabi.load()
    K, n8 = qweight.shape
    N = n8 * 8
    G = K // qzeros.shape[0]
    assert qweight.dtype == torch.int32 and qzeros.dtype == torch.int32 and scales.shape == (K // G, N)
    assert qweight.is_contiguous() and qzeros.is_contiguous() and scales.is_contiguous()
    qweight_t = torch.empty((((N + 255) // 256) * ((K + 127) // 128), 256, 16), dtype=torch.int32, device=qweight.device)
    sz = torch.empty((K // G, N), dtype=torch.int32, device=qweight.device)
    check(lib.b200_awq_repack(ptr(qweight_t), ptr(sz), ptr(qweight), ptr(scales), ptr(qzeros), K, N, G,
                              dtype_code(scales.dtype), stream_ptr(qweight.device)), "b200_awq_repack")
    return qweight_t, sz


def awq_gemm_repacked(x: torch.Tensor, qweight_t: torch.Tensor, sz: torch.Tensor, group_size: int,
                      bias: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Fused dequant + matmul on the streaming layout of `awq_repack`: x [M, K] @ W [K, N] (+bias)."""
    require_cuda(x, qweight_t, sz)
    lib = _cabi.load()
    N = sz.shape[1]
    K = sz.shape[0] * int(group_size)
    assert qweight_t.dtype == torch.int32 and sz.dtype == torch.int32 and qweight_t.is_contiguous() and sz.is_contiguous()
    assert qweight_t.numel() == ((N + 255) // 256) * ((K + 127) // 128) * 4096 and x.shape[-1] == K
    assert x.dtype in (torch.float16, torch.bfloat16)
    x2 = x.reshape(-1, K).contiguous()
    M = x2.shape[0]
    out = _out_buffer(out, M, N, x.dtype, x.device, "awq_gemm")
    if bias is not None:
        bias = bias.to(x.dtype).contiguous()
    ws = _awq_workspace(lib, x.device)
    check(lib.b200_awq_gemm(ptr(out), ptr(x2), ptr(qweight_t), ptr(sz), ptr(bias), M, N, K, int(group_size),
                            dtype_code(x.dtype), ptr(ws), ws.numel(), stream_ptr(x.device)), "b200_awq_gemm")
    return out


def awq_gemm(x: torch.Tensor, qweight: torch.Tensor, scales: torch.Tensor, qzeros: torch.Tensor,
             bias: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Fused dequant + matmul from the CHECKPOINT layout (awq.py:401-418): repacks on every call — for tests and one-off
    use; a layer repacks once (B200AWQLinearMethod.process_weights_after_loading) and calls `awq_gemm_repacked`."""
    assert x.dtype == scales.dtype
    qweight_t, sz = awq_repack(qweight, scales, qzeros)
    return awq_gemm_repacked(x, qweight_t, sz, qweight.shape[0] // qzeros.shape[0], bias=bias, out=out)


# ---- element-wise ----------------------------------------------------------------------------------------
def rmsnorm(x: torch.Tensor, weight: torch.Tensor, eps: float) -> torch.Tensor:
    require_cuda(x, weight)
    lib = _cabi.load()
    H = x.shape[-1]
    x2 = x.reshape(-1, H).contiguous()
    out = torch.empty_like(x2)
    check(lib.b200_rmsnorm(ptr(out), None, ptr(x2), ptr(weight), x2.shape[0], H, float(eps),
                           dtype_code(x.dtype), stream_ptr(x.device)), "b200_rmsnorm")
    return out.view(x.shape)


def fused_add_rmsnorm(x: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor, eps: float) -> None:
    """In place like sgl_kernel.fused_add_rmsnorm: residual += x; x = rmsnorm(residual) * weight."""
    require_cuda(x, residual, weight)
    lib = _cabi.load()
    H = x.shape[-1]
    assert x.is_contiguous() and residual.is_contiguous()
    check(lib.b200_rmsnorm(ptr(x), ptr(residual), ptr(x), ptr(weight), x.numel() // H, H, float(eps),
                           dtype_code(x.dtype), stream_ptr(x.device)), "b200_rmsnorm")


def silu_and_mul(x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    require_cuda(x)
    lib = _cabi.load()
    H = x.shape[-1] // 2
    x2 = x.reshape(-1, 2 * H).contiguous()
    if out is None:
        out = torch.empty(x.shape[:-1] + (H,), device=x.device, dtype=x.dtype)
    check(lib.b200_silu_and_mul(ptr(out), ptr(x2), x2.shape[0], H, dtype_code(x.dtype), stream_ptr(x.device)),
          "b200_silu_and_mul")
    return out


def apply_rope_neox_inplace(positions: torch.Tensor, q: torch.Tensor, k: torch.Tensor, head_dim: int,
                            cos_sin_cache: torch.Tensor) -> None:
    """q [T, Hq*D], k [T, Hkv*D] (may be strided views of a fused qkv), rotated in place."""
    require_cuda(positions, q, k, cos_sin_cache)
    lib = _cabi.load()
    assert positions.dtype == torch.int64 and cos_sin_cache.dtype == torch.float32
    T = q.shape[0]
    hq = q.shape[-1] // head_dim
    hkv = k.shape[-1] // head_dim
    rot = cos_sin_cache.shape[-1]
    assert q.stride(-1) == 1 and k.stride(-1) == 1
    check(lib.b200_rope_neox(ptr(q), ptr(k), ptr(cos_sin_cache), ptr(positions), T, hq, hkv, head_dim, rot,
                             q.stride(0), k.stride(0), dtype_code(q.dtype), stream_ptr(q.device)), "b200_rope_neox")


# ---- fused steps ---------------------------------------------------------------------
class Int8Activation(tuple):
    """(x_q int8 [M, K], x_scale fp32 [M, 1]) produced by a fused producer kernel; accepted by
    B200W8A8Int8LinearMethod.apply in place of the bf16 activation (skips its own quant launch)."""

    __slots__ = ()

    def __new__(cls, x_q: torch.Tensor, x_scale: torch.Tensor, out_dtype: torch.dtype):
        return super().__new__(cls, (x_q, x_scale, out_dtype))

    x_q = property(lambda self: self[0])
    x_scale = property(lambda self: self[1])
    out_dtype = property(lambda self: self[2])

    @property
    def shape(self):
        return self[0].shape


def rmsnorm_quant_int8(x: torch.Tensor, weight: torch.Tensor, eps: float, residual: Optional[torch.Tensor] = None,
                       return_normed: bool = False):
    """(fused_add_)rmsnorm followed by per_token_quant_int8 in ONE launch.  `residual` is updated in place
    (residual += x) like sgl_kernel.fused_add_rmsnorm; x itself is left untouched."""
    require_cuda(x, weight, residual)
    lib = _cabi.load()
    H = x.shape[-1]
    assert x.is_contiguous() and (residual is None or residual.is_contiguous())
    M = x.numel() // H
    x_q = torch.empty((M, H), device=x.device, dtype=torch.int8)
    scale = torch.empty((M, 1), device=x.device, dtype=torch.float32)
    y = torch.empty_like(x) if return_normed else None
    check(lib.b200_rmsnorm_quant_int8(ptr(x_q), ptr(scale), ptr(y), ptr(residual), ptr(x), ptr(weight), M, H, float(eps),
                                      dtype_code(x.dtype), stream_ptr(x.device)), "b200_rmsnorm_quant_int8")
    act = Int8Activation(x_q, scale, x.dtype)
    return (act, y) if return_normed else act


def silu_and_mul_quant_int8(x: torch.Tensor) -> Int8Activation:
    """silu_and_mul followed by per_token_quant_int8 in ONE launch; x is [M, 2H] gate|up."""
    require_cuda(x)
    lib = _cabi.load()
    H = x.shape[-1] // 2
    x2 = x.reshape(-1, 2 * H)
    if not x2.is_contiguous():
        x2 = x2.contiguous()
    M = x2.shape[0]
    x_q = torch.empty((M, H), device=x.device, dtype=torch.int8)
    scale = torch.empty((M, 1), device=x.device, dtype=torch.float32)
    check(lib.b200_silu_and_mul_quant_int8(ptr(x_q), ptr(scale), ptr(x2), M, H, dtype_code(x.dtype), stream_ptr(x.device)),
          "b200_silu_and_mul_quant_int8")
    return Int8Activation(x_q, scale, x.dtype)


class Fp8GroupActivation(tuple):
    """(x_q e4m3 [M, K], x_scale fp32 [M, K / 128], out_dtype): an activation already quantised per (token, 128) group by
    its producer kernel; B200Fp8LinearMethod.apply() takes it in place of the 16-bit activation."""

    def __new__(cls, x_q, x_scale, out_dtype):
        return super().__new__(cls, (x_q, x_scale, out_dtype))

    x_q = property(lambda self: self[0])
    x_scale = property(lambda self: self[1])
    out_dtype = property(lambda self: self[2])

    @property
    def shape(self):
        return self[0].shape


def rmsnorm_group_quant_fp8(x: torch.Tensor, weight: torch.Tensor, eps: float, residual: Optional[torch.Tensor] = None,
                            quant_eps: float = 1e-10) -> Fp8GroupActivation:
    """(fused_add_)rmsnorm followed by sgl_per_token_group_quant_fp8(group 128) in ONE launch; `residual` is updated
    in place (residual += x)."""
    require_cuda(x, weight, residual)
    lib = _cabi.load()
    H = x.shape[-1]
    assert x.is_contiguous() and (residual is None or residual.is_contiguous()) and H % 128 == 0
    M = x.numel() // H
    x_q = torch.empty((M, H), device=x.device, dtype=torch.float8_e4m3fn)
    x_s = torch.empty((M, H // 128), device=x.device, dtype=torch.float32)
    check(lib.b200_rmsnorm_group_quant_fp8(ptr(x_q), ptr(x_s), ptr(residual), ptr(x), ptr(weight), M, H, float(eps),
                                           float(quant_eps), dtype_code(x.dtype), stream_ptr(x.device)),
          "b200_rmsnorm_group_quant_fp8")
    return Fp8GroupActivation(x_q, x_s, x.dtype)


def silu_and_mul_group_quant_fp8(x: torch.Tensor, quant_eps: float = 1e-10) -> Fp8GroupActivation:
    """silu_and_mul followed by sgl_per_token_group_quant_fp8(group 128) in ONE launch; x is [M, 2H] gate|up."""
    require_cuda(x)
    lib = _cabi.load()
    H = x.shape[-1] // 2
    x2 = x.reshape(-1, 2 * H)
    if not x2.is_contiguous():
        x2 = x2.contiguous()
    M = x2.shape[0]
    assert H % 128 == 0
    x_q = torch.empty((M, H), device=x.device, dtype=torch.float8_e4m3fn)
    x_s = torch.empty((M, H // 128), device=x.device, dtype=torch.float32)
    check(lib.b200_silu_and_mul_group_quant_fp8(ptr(x_q), ptr(x_s), ptr(x2), M, H, float(quant_eps), dtype_code(x.dtype),
                                                stream_ptr(x.device)), "b200_silu_and_mul_group_quant_fp8")
    return Fp8GroupActivation(x_q, x_s, x.dtype)


def rope_neox_set_kv(positions: torch.Tensor, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, head_dim: int,
                     cos_sin_cache: torch.Tensor, k_cache: Optional[torch.Tensor], v_cache: Optional[torch.Tensor],
                     loc: Optional[torch.Tensor]) -> None:
    """NeoX RoPE on q / k in place (views of a fused qkv are fine) + set_kv_buffer(k_cache, v_cache, loc, k, v)."""
    require_cuda(positions, q, k, v, cos_sin_cache, k_cache, v_cache, loc)
    lib = _cabi.load()
    assert positions.dtype == torch.int64 and cos_sin_cache.dtype == torch.float32
    T = q.shape[0]
    hq, hkv = q.shape[-1] // head_dim, k.shape[-1] // head_dim
    assert q.stride(-1) == 1 and k.stride(-1) == 1 and v.stride(-1) == 1
    if k_cache is not None:
        assert k_cache.dtype == k.dtype and v_cache.dtype == v.dtype and loc.numel() == T and loc.is_contiguous()
        kc, vc = k_cache.view(k_cache.shape[0], -1), v_cache.view(v_cache.shape[0], -1)
        kcs, vcs, is64 = kc.stride(0), vc.stride(0), _is64(loc)
    else:
        kc = vc = None
        kcs = vcs = is64 = 0
    check(lib.b200_rope_neox_set_kv(ptr(q), ptr(k), ptr(v), ptr(cos_sin_cache), ptr(positions), ptr(kc), ptr(vc), ptr(loc),
                                    is64, T, hq, hkv, head_dim, cos_sin_cache.shape[-1], q.stride(0), k.stride(0),
                                    v.stride(0), kcs, vcs, dtype_code(q.dtype), stream_ptr(q.device)),
          "b200_rope_neox_set_kv")


# ---- caller-side page-table kernels --------------------------------------------------------------------
def _same_int_dtype(*ts: torch.Tensor) -> int:
    dts = {t.dtype for t in ts if t is not None}
    if len(dts) != 1:
        raise RuntimeError(f"index tensors must share one dtype, got {sorted(map(str, dts))}")
    return _is64(next(t for t in ts if t is not None))


def alloc_extend(prefix_lens: torch.Tensor, seq_lens: torch.Tensor, last_loc: Optional[torch.Tensor],
                 free_pages: torch.Tensor, out_indices: torch.Tensor, ret_values: torch.Tensor, page_size: int,
                 req_to_token: Optional[torch.Tensor] = None, req_pool_indices: Optional[torch.Tensor] = None,
                 write_req_to_token: bool = False) -> None:
    """alloc_extend_kernel (mem_cache/allocator.py:276-360).  last_loc=None reads it from req_to_token (get_last_loc
    fused); write_req_to_token=True also stores the new slots into req_to_token[req, prefix:seq]."""
    require_cuda(prefix_lens, seq_lens, last_loc, free_pages, out_indices, ret_values, req_to_token, req_pool_indices)
    lib = _cabi.load()
    assert free_pages.dtype == torch.int64 and out_indices.dtype == torch.int64 and ret_values.dtype == torch.int64
    assert req_to_token is None or (req_to_token.dtype == torch.int32 and req_to_token.stride(1) == 1)
    check(lib.b200_alloc_extend(ptr(out_indices), ptr(ret_values), ptr(prefix_lens), ptr(seq_lens), ptr(last_loc),
                                _same_int_dtype(prefix_lens, seq_lens, last_loc), ptr(free_pages), seq_lens.numel(),
                                int(page_size), ptr(req_to_token), 0 if req_to_token is None else req_to_token.stride(0),
                                ptr(req_pool_indices), 0 if req_pool_indices is None else _is64(req_pool_indices),
                                int(write_req_to_token), stream_ptr(seq_lens.device)), "b200_alloc_extend")


def alloc_decode(seq_lens: torch.Tensor, last_loc: Optional[torch.Tensor], free_pages: torch.Tensor,
                 out_indices: torch.Tensor, ret_values: torch.Tensor, page_size: int,
                 req_to_token: Optional[torch.Tensor] = None, req_pool_indices: Optional[torch.Tensor] = None,
                 write_req_to_token: bool = False) -> None:
    """alloc_decode_kernel (allocator.py:364-401), same fusions as alloc_extend."""
    require_cuda(seq_lens, last_loc, free_pages, out_indices, ret_values, req_to_token, req_pool_indices)
    lib = _cabi.load()
    assert free_pages.dtype == torch.int64 and out_indices.dtype == torch.int64 and ret_values.dtype == torch.int64
    assert req_to_token is None or (req_to_token.dtype == torch.int32 and req_to_token.stride(1) == 1)
    check(lib.b200_alloc_decode(ptr(out_indices), ptr(ret_values), ptr(seq_lens), ptr(last_loc),
                                _same_int_dtype(seq_lens, last_loc), ptr(free_pages), seq_lens.numel(), int(page_size),
                                ptr(req_to_token), 0 if req_to_token is None else req_to_token.stride(0),
                                ptr(req_pool_indices), 0 if req_pool_indices is None else _is64(req_pool_indices),
                                int(write_req_to_token), stream_ptr(seq_lens.device)), "b200_alloc_decode")


def write_req_to_token_pool(req_to_token: torch.Tensor, req_pool_indices: torch.Tensor, prefix_lens: torch.Tensor,
                            seq_lens: torch.Tensor, extend_lens: torch.Tensor, out_cache_loc: torch.Tensor) -> None:
    """write_req_to_token_pool_triton (managers/schedule_batch.py:1920-1953)."""
    require_cuda(req_to_token, req_pool_indices, prefix_lens, seq_lens, extend_lens, out_cache_loc)
    lib = _cabi.load()
    assert req_to_token.dtype == torch.int32 and req_to_token.stride(1) == 1
    check(lib.b200_write_req_to_token_pool(ptr(req_to_token), req_to_token.stride(0), ptr(req_pool_indices),
                                           _is64(req_pool_indices), ptr(prefix_lens), ptr(seq_lens), ptr(extend_lens),
                                           _same_int_dtype(prefix_lens, seq_lens, extend_lens), ptr(out_cache_loc),
                                           _is64(out_cache_loc), req_pool_indices.numel(),
                                           stream_ptr(req_to_token.device)), "b200_write_req_to_token_pool")


def get_last_loc(req_to_token: torch.Tensor, req_pool_indices: torch.Tensor, prefix_lens: torch.Tensor) -> torch.Tensor:
    """get_last_loc (schedule_batch.py:1956-2027): result has prefix_lens' dtype."""
    require_cuda(req_to_token, req_pool_indices, prefix_lens)
    lib = _cabi.load()
    assert req_to_token.dtype == torch.int32 and req_to_token.stride(1) == 1
    result = torch.empty_like(prefix_lens)
    check(lib.b200_get_last_loc(ptr(result), ptr(req_to_token), req_to_token.stride(0), ptr(req_pool_indices),
                                _is64(req_pool_indices), ptr(prefix_lens), _is64(prefix_lens), prefix_lens.numel(),
                                stream_ptr(prefix_lens.device)), "b200_get_last_loc")
    return result


def compute_position(extend_prefix_lens: Optional[torch.Tensor], extend_seq_lens: torch.Tensor,
                     extend_seq_lens_sum: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """compute_position_triton (model_executor/forward_batch_info.py:886-910): (positions int64, extend_start_loc int32)."""
    require_cuda(extend_prefix_lens, extend_seq_lens)
    lib = _cabi.load()
    bs = extend_seq_lens.numel()
    if extend_prefix_lens is not None and extend_prefix_lens.numel() != bs:
        extend_prefix_lens = None  # the reference's has_prefix rule (:890)
    positions = torch.empty(extend_seq_lens_sum, dtype=torch.int64, device=extend_seq_lens.device)
    start = torch.empty(bs, dtype=torch.int32, device=extend_seq_lens.device)
    check(lib.b200_compute_position(ptr(positions), ptr(start), ptr(extend_prefix_lens), ptr(extend_seq_lens),
                                    _same_int_dtype(extend_prefix_lens, extend_seq_lens), bs,
                                    stream_ptr(extend_seq_lens.device)), "b200_compute_position")
    return positions, start


def create_flashinfer_kv_indices(req_to_token: torch.Tensor, req_pool_indices: torch.Tensor,
                                 page_kernel_lens: torch.Tensor, kv_indptr: torch.Tensor,
                                 kv_start_idx: Optional[torch.Tensor], kv_indices: torch.Tensor) -> None:
    """create_flashinfer_kv_indices_triton (layers/attention/utils.py:11-47)."""
    require_cuda(req_to_token, req_pool_indices, page_kernel_lens, kv_indptr, kv_start_idx, kv_indices)
    lib = _cabi.load()
    assert req_to_token.dtype == torch.int32 and req_to_token.stride(1) == 1
    assert kv_indptr.dtype == torch.int32 and kv_indices.dtype == torch.int32
    assert kv_start_idx is None or kv_start_idx.dtype == torch.int32
    check(lib.b200_create_kv_indices(ptr(kv_indices), ptr(req_to_token), req_to_token.stride(0), ptr(req_pool_indices),
                                     _is64(req_pool_indices), ptr(page_kernel_lens), _is64(page_kernel_lens),
                                     ptr(kv_indptr), ptr(kv_start_idx), req_pool_indices.numel(),
                                     stream_ptr(req_to_token.device)), "b200_create_kv_indices")


# ---- greedy sampling ------------------------------------------------------------------------------------------------------
_ARGMAX_WS = {}


def argmax(logits: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """torch.argmax(logits, -1) for [rows, V] bf16/fp16 logits (lowest index on ties), rows split over CTAs."""
    require_cuda(logits, out)
    lib = _cabi.load()
    assert logits.dim() == 2 and logits.stride(1) == 1
    rows, V = logits.shape
    if out is None:
        out = torch.empty(rows, dtype=torch.int64, device=logits.device)
    assert out.dtype == torch.int64 and out.is_contiguous() and out.numel() == rows
    ws = _ARGMAX_WS.get(logits.device.index)
    if ws is None or ws.numel() < rows * 16:
        ws = torch.zeros(max(rows, 1024) * 16, dtype=torch.uint8, device=logits.device)
        _ARGMAX_WS[logits.device.index] = ws
    check(lib.b200_argmax(ptr(out), ptr(logits), rows, V, logits.stride(0), dtype_code(logits.dtype), ptr(ws), ws.numel(),
                          stream_ptr(logits.device)), "b200_argmax")
    return out


def argmax_key(logits: torch.Tensor, index_offset: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Per-row int64 key (value bits, ~global index) of this rank's [rows, V_local] logits slice: max-combine the
    keys over the TP ranks, then `argmax_key_decode`.  Vocab-parallel replacement of all-gather + argmax."""
    require_cuda(logits, out)
    lib = _cabi.load()
    assert logits.dim() == 2 and logits.stride(1) == 1
    rows, V = logits.shape
    if out is None:
        out = torch.empty(rows, dtype=torch.int64, device=logits.device)
    assert out.dtype == torch.int64 and out.is_contiguous() and out.numel() == rows
    ws = _ARGMAX_WS.get(logits.device.index)
    if ws is None or ws.numel() < rows * 16:
        ws = torch.zeros(max(rows, 1024) * 16, dtype=torch.uint8, device=logits.device)
        _ARGMAX_WS[logits.device.index] = ws
    check(lib.b200_argmax_key(ptr(out), ptr(logits), rows, V, logits.stride(0), dtype_code(logits.dtype), int(index_offset),
                              ptr(ws), ws.numel(), stream_ptr(logits.device)), "b200_argmax_key")
    return out


def argmax_key_decode(keys: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    require_cuda(keys, out)
    lib = _cabi.load()
    assert keys.dtype == torch.int64 and keys.is_contiguous()
    if out is None:
        out = torch.empty_like(keys)
    check(lib.b200_argmax_key_decode(ptr(out), ptr(keys), keys.numel(), stream_ptr(keys.device)), "b200_argmax_key_decode")
    return out


def embedding(ids: torch.Tensor, table: torch.Tensor, seq_lens: Optional[torch.Tensor] = None,
              positions: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """table[ids] (one launch); with seq_lens + positions also writes positions = seq_lens - 1 (decode prologue)."""
    require_cuda(ids, table, seq_lens, positions, out)
    lib = _cabi.load()
    assert ids.dtype == torch.int64 and ids.is_contiguous() and table.is_contiguous() and table.dim() == 2
    T, (V, H) = ids.numel(), table.shape
    if out is None:
        out = torch.empty((T, H), dtype=table.dtype, device=table.device)
    if seq_lens is not None:
        assert seq_lens.dtype == torch.int64 and positions.dtype == torch.int64 and positions.numel() == T
    check(lib.b200_embedding(ptr(out), ptr(table), ptr(ids), T, H, V, dtype_code(table.dtype), ptr(positions), ptr(seq_lens),
                             stream_ptr(table.device)), "b200_embedding")
    return out
