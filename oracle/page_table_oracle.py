"""TEST INFRASTRUCTURE ONLY — never imported by the product path.

CPU restatement (plain Python / torch integer loops) of the reference's caller-side page-table
kernels, SURVEY §8f #1.  Every function cites the reference code it follows (paths under
/root/reference/python/sglang/srt/).  Pinned by tests/golden/page_table_*.pt, which
tests/golden/make_golden_page_table.py produced by running the reference's own torch
implementations (alloc_extend_native, alloc_decode_kernel_ascend, compute_position_torch,
imported from the unmodified tree) plus a line-by-line emulation of the Triton programs.
"""
from typing import List, Optional, Tuple

import torch


def _ceil_div(a: int, b: int) -> int:
    return (a + b - 1) // b


def alloc_extend(prefix_lens, seq_lens, last_loc, free_pages, page_size: int) -> Tuple[torch.Tensor, int, int]:
    """alloc_extend_kernel (mem_cache/allocator.py:276-360).  Returns (out_indices int64, new_pages, sum_extend);
    the kernel packs the last two as (new_pages << 32) | sum_extend (:312-317)."""
    P = page_size
    out: List[int] = []
    page_ptr = 0
    for pre, seq, ll in zip(prefix_lens.tolist(), seq_lens.tolist(), last_loc.tolist()):
        n_new = _ceil_div(seq, P) - _ceil_div(pre, P)
        # part 1: rest of the old partial page (:319-330)
        n1 = min(seq, _ceil_div(pre, P) * P) - pre
        out += [ll + 1 + j for j in range(n1)]
        # part 2: whole new pages (:334-350)
        n2 = max(seq // P * P - _ceil_div(pre, P) * P, 0) if pre + n1 < seq else 0
        for j in range(n2):
            out.append(int(free_pages[page_ptr + j // P]) * P + j % P)
        # part 3: the new partial page = the last new page of this request (:352-360)
        if pre + n1 + n2 < seq:
            n3 = seq - seq // P * P
            start = int(free_pages[page_ptr + n_new - 1])
            out += [start * P + j for j in range(n3)]
        page_ptr += n_new
    return torch.tensor(out, dtype=torch.int64), page_ptr, len(out)


def alloc_decode(seq_lens, last_loc, free_pages, page_size: int) -> Tuple[torch.Tensor, int]:
    """alloc_decode_kernel (allocator.py:364-401): a request whose new token opens a page takes the next free page,
    otherwise last_loc + 1."""
    P = page_size
    out = []
    page_ptr = 0
    for seq, ll in zip(seq_lens.tolist(), last_loc.tolist()):
        n_new = _ceil_div(seq, P) - _ceil_div(seq - 1, P)
        if n_new == 0:
            out.append(ll + 1)
        else:
            out.append(int(free_pages[page_ptr]) * P)
        page_ptr += n_new
    return torch.tensor(out, dtype=torch.int64), page_ptr


def get_last_loc(req_to_token, req_pool_indices, prefix_lens) -> torch.Tensor:
    """get_last_loc_torch (managers/schedule_batch.py:1972-1982)."""
    return torch.where(prefix_lens > 0, req_to_token[req_pool_indices, (prefix_lens - 1).clamp(min=0)].to(prefix_lens.dtype),
                       torch.full_like(prefix_lens, -1))


def write_req_to_token_pool(req_to_token, req_pool_indices, prefix_lens, seq_lens, out_cache_loc) -> None:
    """The loop the reference falls back to next to write_req_to_token_pool_triton (schedule_batch.py:1294-1316):
    req_to_token[req, pre:seq] = out_cache_loc[pt : pt + seq - pre]."""
    pt = 0
    for r, pre, seq in zip(req_pool_indices.tolist(), prefix_lens.tolist(), seq_lens.tolist()):
        n = seq - pre
        req_to_token[r, pre:seq] = out_cache_loc[pt:pt + n].to(req_to_token.dtype)
        pt += n


def compute_position(extend_prefix_lens, extend_seq_lens) -> Tuple[torch.Tensor, torch.Tensor]:
    """compute_position_torch (model_executor/forward_batch_info.py:942-957)."""
    pos = [torch.arange(p, p + e, dtype=torch.int64) for p, e in zip(extend_prefix_lens.tolist(), extend_seq_lens.tolist())]
    positions = torch.cat(pos) if pos else torch.zeros(0, dtype=torch.int64)
    start = torch.zeros_like(extend_seq_lens)
    start[1:] = torch.cumsum(extend_seq_lens[:-1], dim=0)
    return positions, start.to(torch.int32)


def create_kv_indices(req_to_token, req_pool_indices, page_kernel_lens, kv_indptr, kv_start_idx: Optional[torch.Tensor],
                      total: int) -> torch.Tensor:
    """create_flashinfer_kv_indices_triton (layers/attention/utils.py:11-47)."""
    out = torch.zeros(total, dtype=torch.int32)
    for b, r in enumerate(req_pool_indices.tolist()):
        s = int(kv_start_idx[b]) if kv_start_idx is not None else 0
        n = int(page_kernel_lens[b])
        o = int(kv_indptr[b])
        out[o:o + n] = req_to_token[r, s:s + n]
    return out
