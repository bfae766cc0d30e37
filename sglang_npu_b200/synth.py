"""Deterministic synthetic inputs for the hot path (SURVEY §8d "Synthetic inputs").

Used by bench.py, the tests and the golden-vector generator so that the oracle and the device
path see identical data.  Page tables follow the paged allocator's contract
(reference python/sglang/srt/mem_cache/allocator.py:428-542): slots 0..page_size-1 are the
reserved padding page, token t of a request lives at offset t % page_size inside its page, and
a request's pages are NOT physically contiguous (pages are dealt from a random permutation).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Sequence

import torch


@dataclass
class PagedBatch:
    req_to_token: torch.Tensor      # [max_reqs, width] int32
    req_pool_indices: torch.Tensor  # [bs] int64
    seq_lens: torch.Tensor          # [bs] int64
    num_slots: int                  # KV pool `size` (excluding the padding page)
    page_size: int


def make_paged_batch(seq_lens: Sequence[int], page_size: int, *, seed: int = 1234, width: Optional[int] = None,
                     max_reqs: Optional[int] = None, shared_prefix: int = 0, extra_pages: int = 3,
                     reserve_tokens: int = 0) -> PagedBatch:
    """Build a page table for requests of the given lengths (CPU tensors; move with .to()).

    shared_prefix: the first `shared_prefix` tokens (multiple of page_size) of every request alias
    the same physical pages (radix-cache prefix reuse, cfg 3 / cfg 5).
    reserve_tokens: extra per-request capacity (e.g. decode steps to come) that also gets pages.
    """
    g = torch.Generator().manual_seed(seed)
    bs = len(seq_lens)
    P = page_size
    assert shared_prefix % P == 0
    cap = [int(n) + reserve_tokens for n in seq_lens]
    pages_per_req = [(max(c - shared_prefix, 0) + P - 1) // P for c in cap]
    shared_pages = shared_prefix // P
    total_pages = sum(pages_per_req) + shared_pages + extra_pages
    perm = (torch.randperm(total_pages, generator=g) + 1).tolist()  # page 0 is the padding page
    width = width or (max(cap) + 4)
    max_reqs = max_reqs or (bs + 3)
    r2t = torch.zeros((max_reqs, width), dtype=torch.int32)
    rows = torch.randperm(max_reqs, generator=g)[:bs]
    cursor = 0
    shared = perm[cursor : cursor + shared_pages]
    cursor += shared_pages
    offs = torch.arange(P, dtype=torch.int32)
    for i in range(bs):
        pages = list(shared) + perm[cursor : cursor + pages_per_req[i]]
        cursor += pages_per_req[i]
        if pages:
            slots = (torch.tensor(pages, dtype=torch.int32).view(-1, 1) * P + offs.view(1, -1)).reshape(-1)
            n = min(slots.numel(), width)
            r2t[rows[i], :n] = slots[:n]
    return PagedBatch(req_to_token=r2t, req_pool_indices=rows.to(torch.int64),
                      seq_lens=torch.tensor(list(seq_lens), dtype=torch.int64),
                      num_slots=(total_pages + 1) * P - P, page_size=P)


def randn_kv(num_slots: int, page_size: int, num_kv_heads: int, head_dim: int, dtype=torch.bfloat16,
             seed: int = 1234, device="cpu"):
    """K and V pools [num_slots + page_size, Hkv, D] filled N(0,1) (test/srt/cpu/test_decode.py:73-80)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    shape = (num_slots + page_size, num_kv_heads, head_dim)
    k = torch.randn(shape, generator=g, dtype=torch.float32).to(dtype)
    v = torch.randn(shape, generator=g, dtype=torch.float32).to(dtype)
    return k.to(device), v.to(device)


def randn_kv_device(num_slots: int, page_size: int, num_kv_heads: int, head_dim: int, dtype, device, seed=1234):
    """Same shape, generated on the device (full-size bench pools are too big to ship from host)."""
    g = torch.Generator(device=device).manual_seed(seed)
    shape = (num_slots + page_size, num_kv_heads, head_dim)
    k = torch.empty(shape, dtype=dtype, device=device).normal_(generator=g)
    v = torch.empty(shape, dtype=dtype, device=device).normal_(generator=g)
    return k, v


def int8_weight(n: int, k: int, seed: int = 1234):
    """w_q ~ U{-127..127} int8 [N,K], w_s ~ U(0, 1e-2) fp32 [N,1] (test/srt/cpu/test_gemm.py:92-94)."""
    g = torch.Generator().manual_seed(seed)
    w = torch.randint(-127, 128, (n, k), generator=g, dtype=torch.int32).to(torch.int8)
    s = torch.rand((n, 1), generator=g, dtype=torch.float32) * 1e-2
    return w, s


def fp8_block_weight(n: int, k: int, seed: int = 1234, block=(128, 128)):
    """e4m3 weights U(-1,1)*448 and |N(0,1)|*1e-3 block scales (sgl-kernel/tests/test_fp8_blockwise_gemm.py:66-75)."""
    g = torch.Generator().manual_seed(seed)
    w = ((torch.rand((n, k), generator=g) - 0.5) * 2 * 448.0).clamp(-448, 448).to(torch.float8_e4m3fn)
    s = torch.randn(((n + block[0] - 1) // block[0], (k + block[1] - 1) // block[1]), generator=g).abs() * 1e-3
    return w, s.float()


def awq_weight(k: int, n: int, group_size: int = 128, dtype=torch.float16, seed: int = 1234):
    """qweight/qzeros ~ randint(0, 2^31-1) int32, scales ~ U(0,1)/sqrt(K)-ish (test_awq_dequant.py:81-101)."""
    g = torch.Generator().manual_seed(seed)
    qweight = torch.randint(0, 2**31 - 1, (k, n // 8), generator=g, dtype=torch.int64).to(torch.int32)
    qzeros = torch.randint(0, 2**31 - 1, (k // group_size, n // 8), generator=g, dtype=torch.int64).to(torch.int32)
    scales = (torch.rand((k // group_size, n), generator=g) * 0.02).to(dtype)
    return qweight, scales, qzeros


def activations(m: int, k: int, dtype=torch.bfloat16, seed: int = 1234, scale: float = 0.1):
    """x ~ N(0,1)/10 (test/srt/cpu/test_gemm.py:85)."""
    g = torch.Generator().manual_seed(seed)
    return (torch.randn((m, k), generator=g) * scale).to(dtype)
