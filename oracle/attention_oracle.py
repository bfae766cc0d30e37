"""TEST INFRASTRUCTURE (oracle) — CPU restatement of the reference attention path.

Parity: PINNED.  This restatement is checked bit-for-bit / to fp tolerance against the
UNMODIFIED reference classes (TorchNativeAttnBackend, MHATokenToKVPool, ReqToTokenPool,
RadixAttention, ForwardBatch) imported from /root/reference via oracle/ref_shims.py, and the
outputs of that real reference are committed as golden vectors (tests/golden/attn_*.pt, made by
tests/golden/make_golden.py); tests/test_oracle_golden.py re-checks the restatement against
those vectors wherever the tests run.

Each function cites the reference lines it follows.  Everything here is plain torch on CPU.
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch
import torch.nn.functional as F


def set_kv_buffer(k_cache: torch.Tensor, v_cache: torch.Tensor, loc: torch.Tensor,
                  cache_k: torch.Tensor, cache_v: torch.Tensor) -> None:
    """k_buffer[layer][loc] = cache_k; v_buffer[layer][loc] = cache_v
    (python/sglang/srt/mem_cache/memory_pool.py:405-407)."""
    k_cache[loc] = cache_k
    v_cache[loc] = cache_v


def set_kv_buffer_fp8(k_cache_u8: torch.Tensor, v_cache_u8: torch.Tensor, loc: torch.Tensor,
                      cache_k: torch.Tensor, cache_v: torch.Tensor, k_scale: Optional[float] = None,
                      v_scale: Optional[float] = None) -> None:
    """The cast branch of MHATokenToKVPool.set_kv_buffer for an fp8_e4m3 pool (memory_pool.py:385-395): optional
    in-place division by the scale (done on a copy here), `.to(float8_e4m3fn)`, stored through the uint8 view
    (memory_pool.py:114-118)."""
    ck, cv = cache_k.clone(), cache_v.clone()
    if k_scale is not None:
        ck.div_(k_scale)
    if v_scale is not None:
        cv.div_(v_scale)
    k_cache_u8[loc] = ck.to(torch.float8_e4m3fn).view(torch.uint8)
    v_cache_u8[loc] = cv.to(torch.float8_e4m3fn).view(torch.uint8)


def decode_attention_fp8kv(q: torch.Tensor, k_cache_u8: torch.Tensor, v_cache_u8: torch.Tensor,
                           req_to_token: torch.Tensor, req_pool_indices: torch.Tensor, seq_lens: torch.Tensor,
                           scaling: float, k_scale: Optional[float] = None, v_scale: Optional[float] = None,
                           compute_dtype: Optional[torch.dtype] = torch.float32) -> torch.Tensor:
    """Decode over an fp8_e4m3 cache.  torch_native has no fp8 path (SDPA rejects mixed dtypes), so this restates the
    reference's Triton kernel: K is up-cast to the q dtype before the dot (triton_ops/decode_attention.py:336,
    exact for e4m3 -> bf16/f16), V likewise here (the Triton kernel instead rounds P to e4m3, :373 — a precision
    loss this build does not copy).  k_scale / v_scale undo set_kv_buffer's division (FlashInfer-style read side);
    None = the Triton backend's behaviour."""
    k = k_cache_u8.view(torch.float8_e4m3fn).to(torch.float32) * (1.0 if k_scale is None else float(k_scale))
    v = v_cache_u8.view(torch.float8_e4m3fn).to(torch.float32) * (1.0 if v_scale is None else float(v_scale))
    return decode_attention(q, k, v, req_to_token, req_pool_indices, seq_lens, scaling,
                            compute_dtype=compute_dtype).to(q.dtype)


def extend_attention_fp8kv(q: torch.Tensor, k_new: torch.Tensor, v_new: torch.Tensor, k_cache_u8: torch.Tensor,
                           v_cache_u8: torch.Tensor, req_to_token: torch.Tensor, req_pool_indices: torch.Tensor,
                           seq_lens: torch.Tensor, extend_prefix_lens: torch.Tensor, extend_seq_lens: torch.Tensor,
                           out_cache_loc: torch.Tensor, scaling: float, k_scale: Optional[float] = None,
                           v_scale: Optional[float] = None, causal: bool = True) -> torch.Tensor:
    """Extend over an fp8_e4m3 cache the way the reference's Triton extend kernel sees it
    (triton_ops/extend_attention.py:302-440): prefix keys/values are the cache's e4m3 values up-cast (times the scale),
    the new tokens are the un-quantised activations (k_extend / v_extend).  Evaluated with the torch_native extend
    formula in fp32 on a temporary float cache holding exactly those values."""
    kf = k_cache_u8.view(torch.float8_e4m3fn).to(torch.float32) * (1.0 if k_scale is None else float(k_scale))
    vf = v_cache_u8.view(torch.float8_e4m3fn).to(torch.float32) * (1.0 if v_scale is None else float(v_scale))
    kf[out_cache_loc] = k_new.to(torch.float32)
    vf[out_cache_loc] = v_new.to(torch.float32)
    return extend_attention(q, kf, vf, req_to_token, req_pool_indices, seq_lens, extend_prefix_lens, extend_seq_lens,
                            scaling, causal=causal, compute_dtype=torch.float32).to(q.dtype)


def _sdpa(q, k, v, scaling, causal, enable_gqa):
    # q [H, Lq, D], k/v [Hkv, Lk, D] — torch_native_backend.py:97-107 / :164-174
    return F.scaled_dot_product_attention(q.unsqueeze(0), k.unsqueeze(0), v.unsqueeze(0),
                                          enable_gqa=enable_gqa, scale=scaling, is_causal=causal).squeeze(0)


def decode_attention(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                     req_to_token: torch.Tensor, req_pool_indices: torch.Tensor, seq_lens: torch.Tensor,
                     scaling: float, compute_dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    """TorchNativeAttnBackend._run_sdpa_forward_decode (torch_native_backend.py:112-180).

    q [bs, Hq, D]; caches [slots, Hkv, D]; returns o [bs, Hq, D].  `compute_dtype=torch.float32`
    evaluates the same formula in fp32 (a tighter yardstick than the bf16 reference run).
    """
    bs, hq, d = q.shape
    hkv = k_cache.shape[1]
    out = torch.empty_like(q)
    qh = q.movedim(0, 1)  # [Hq, bs, D]
    for i in range(bs):
        n = int(seq_lens[i])
        if n == 0:
            out[i] = 0
            continue
        toks = req_to_token[int(req_pool_indices[i]), :n].long()
        k = k_cache[toks].movedim(0, 1)
        v = v_cache[toks].movedim(0, 1)
        qi = qh[:, i : i + 1, :]
        if compute_dtype is not None:
            qi, k, v = qi.to(compute_dtype), k.to(compute_dtype), v.to(compute_dtype)
        o = _sdpa(qi, k, v, scaling, False, hq != hkv)
        out[i] = o.movedim(1, 0)[0].to(out.dtype)
    return out


def extend_attention(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                     req_to_token: torch.Tensor, req_pool_indices: torch.Tensor, seq_lens: torch.Tensor,
                     extend_prefix_lens: torch.Tensor, extend_seq_lens: torch.Tensor, scaling: float,
                     causal: bool = True, compute_dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    """TorchNativeAttnBackend._run_sdpa_forward_extend (torch_native_backend.py:27-110).

    The reference pads the query to seq_len rows (rows < prefix are uninitialised and discarded,
    :81-87); zeros are used here for those rows — they do not influence the kept rows.
    """
    T, hq, d = q.shape
    hkv = k_cache.shape[1]
    out = torch.empty_like(q)
    qh = q.movedim(0, 1)
    start = 0
    for i in range(seq_lens.shape[0]):
        e = int(extend_seq_lens[i])
        p = int(extend_prefix_lens[i])
        n = int(seq_lens[i])
        toks = req_to_token[int(req_pool_indices[i]), :n].long()
        k = k_cache[toks].movedim(0, 1)
        v = v_cache[toks].movedim(0, 1)
        q_red = torch.zeros((hq, n, d), dtype=q.dtype)
        q_red[:, p:, :] = qh[:, start : start + e, :]
        if compute_dtype is not None:
            q_red, k, v = q_red.to(compute_dtype), k.to(compute_dtype), v.to(compute_dtype)
        o = _sdpa(q_red, k, v, scaling, causal, hq != hkv)
        out[start : start + e] = o.movedim(1, 0)[p:].to(out.dtype)
        start += e
    return out


def merge_state(v_a: torch.Tensor, s_a: torch.Tensor, v_b: torch.Tensor, s_b: torch.Tensor
                ) -> Tuple[torch.Tensor, torch.Tensor]:
    """merge_attn_states_kernel (sgl-kernel/csrc/attention/merge_attn_states.cu:31-110); same formula
    as the reference test's torch restatement (sgl-kernel/tests/test_merge_state_v2.py)."""
    p = torch.where(torch.isinf(s_a), torch.full_like(s_a, -math.inf), s_a.float())
    s = torch.where(torch.isinf(s_b), torch.full_like(s_b, -math.inf), s_b.float())
    m = torch.maximum(p, s)
    pe = torch.exp(p - m)
    se = torch.exp(s - m)
    tot = pe + se
    out = v_a.float() * (pe / tot).unsqueeze(-1) + v_b.float() * (se / tot).unsqueeze(-1)
    return out.to(v_a.dtype), torch.log(tot) + m
