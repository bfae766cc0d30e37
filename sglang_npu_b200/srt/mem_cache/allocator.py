"""Paged KV-slot allocator driving the page-table kernels — the host-side mirror of the reference's
`PagedTokenToKVPoolAllocator` (python/sglang/srt/mem_cache/allocator.py:404-571): same constructor state
(`free_pages` int64 starting at page 1, `release_pages`), same `alloc / alloc_extend / alloc_decode / free /
merge_and_sort_free / clear` contract and return values, so the reference scheduler can drive it unchanged.

What differs is underneath: the Triton `alloc_extend_kernel` / `alloc_decode_kernel` are the CUDA kernels of
csrc/page_table.cu, and the two fusions the reference leaves as a TODO (allocator.py:411) are exposed as
`alloc_extend_fused / alloc_decode_fused`: last_loc is read from `req_to_token` and the new slots are written
back to it in the same launch (get_last_loc + alloc + write_req_to_token_pool -> one launch).
"""
from __future__ import annotations

from typing import Optional

import torch

from ... import kernels


class PagedTokenToKVPoolAllocator:
    def __init__(self, size: int, page_size: int, dtype: torch.dtype, device, kvcache=None):
        self.size = size
        self.page_size = page_size
        self.dtype = dtype
        self.device = device
        self._kvcache = kvcache
        self.num_pages = size // page_size
        self.ret_values = torch.empty((), dtype=torch.int64, device=device)
        self.clear()

    # ---- bookkeeping (allocator.py:52-107, :555-571) ---------------------------------------------------------------
    def clear(self):
        # page 0 is the padding page that padded tokens write to
        self.free_pages = torch.arange(1, self.num_pages + 1, dtype=torch.int64, device=self.device)
        self.release_pages = torch.empty((0,), dtype=torch.int64, device=self.device)
        self.is_not_in_free_group = True
        self.free_group = []

    def available_size(self) -> int:
        return (len(self.free_pages) + len(self.release_pages)) * self.page_size

    def merge_and_sort_free(self):
        if len(self.release_pages) > 0:
            self.free_pages = torch.cat((self.free_pages, self.release_pages))
            self.free_pages, _ = torch.sort(self.free_pages)
            self.release_pages = torch.empty((0,), dtype=self.release_pages.dtype, device=self.device)

    def free(self, free_index: torch.Tensor):
        if free_index.numel() == 0:
            return
        if self.is_not_in_free_group:
            free_page_indices = torch.unique(free_index // self.page_size)
            self.release_pages = torch.cat((free_page_indices, self.release_pages))
        else:
            self.free_group.append(free_index)

    def free_group_begin(self):
        self.is_not_in_free_group = False
        self.free_group = []

    def free_group_end(self):
        self.is_not_in_free_group = True
        if self.free_group:
            self.free(torch.cat(self.free_group))

    # ---- allocation ---------------------------------------------------------------------------------------------------
    def alloc(self, need_size: int) -> Optional[torch.Tensor]:
        """Page-aligned allocation of whole pages (allocator.py:428-449)."""
        num_pages = need_size // self.page_size
        if num_pages > len(self.free_pages):
            self.merge_and_sort_free()
        if num_pages > len(self.free_pages):
            return None
        out_pages = self.free_pages[:num_pages]
        self.free_pages = self.free_pages[num_pages:]
        return (out_pages[:, None] * self.page_size + torch.arange(self.page_size, device=self.device)).reshape(-1)

    def _ensure_pages(self, prefix_lens: torch.Tensor, seq_lens: torch.Tensor) -> bool:
        P = self.page_size
        est = int((((seq_lens + P - 1) // P) - ((prefix_lens + P - 1) // P)).sum().item())
        if est > len(self.free_pages):
            self.merge_and_sort_free()
        return est <= len(self.free_pages)

    def alloc_extend(self, prefix_lens: torch.Tensor, seq_lens: torch.Tensor, last_loc: torch.Tensor,
                     extend_num_tokens: int) -> Optional[torch.Tensor]:
        """allocator.py:451-503."""
        self._ensure_pages(prefix_lens, seq_lens)
        out_indices = torch.empty((extend_num_tokens,), dtype=torch.int64, device=self.device)
        kernels.alloc_extend(prefix_lens, seq_lens, last_loc, self.free_pages, out_indices, self.ret_values,
                             self.page_size)
        return self._commit(int(self.ret_values.item()) >> 32, out_indices)

    def alloc_decode(self, seq_lens: torch.Tensor, last_loc: torch.Tensor) -> Optional[torch.Tensor]:
        """allocator.py:505-542."""
        self._ensure_pages(seq_lens - 1, seq_lens)
        out_indices = torch.empty((seq_lens.numel(),), dtype=torch.int64, device=self.device)
        kernels.alloc_decode(seq_lens, last_loc, self.free_pages, out_indices, self.ret_values, self.page_size)
        return self._commit(int(self.ret_values.item()), out_indices)

    def alloc_extend_fused(self, req_to_token: torch.Tensor, req_pool_indices: torch.Tensor, prefix_lens: torch.Tensor,
                           seq_lens: torch.Tensor, extend_num_tokens: int) -> Optional[torch.Tensor]:
        """get_last_loc + alloc_extend + write_req_to_token_pool in one launch.  Unlike the un-fused call it must not
        run when the pool is short of pages (it would write slots of pages it does not own into req_to_token)."""
        if not self._ensure_pages(prefix_lens, seq_lens):
            return None
        out_indices = torch.empty((extend_num_tokens,), dtype=torch.int64, device=self.device)
        kernels.alloc_extend(prefix_lens, seq_lens, None, self.free_pages, out_indices, self.ret_values, self.page_size,
                             req_to_token=req_to_token, req_pool_indices=req_pool_indices, write_req_to_token=True)
        return self._commit(int(self.ret_values.item()) >> 32, out_indices)

    def alloc_decode_fused(self, req_to_token: torch.Tensor, req_pool_indices: torch.Tensor,
                           seq_lens: torch.Tensor) -> Optional[torch.Tensor]:
        # a decode step needs at most one new page per request: skip the device-side estimate (and its host sync)
        # whenever the free list covers that bound
        if seq_lens.numel() > len(self.free_pages) and not self._ensure_pages(seq_lens - 1, seq_lens):
            return None
        out_indices = torch.empty((seq_lens.numel(),), dtype=torch.int64, device=self.device)
        kernels.alloc_decode(seq_lens, None, self.free_pages, out_indices, self.ret_values, self.page_size,
                             req_to_token=req_to_token, req_pool_indices=req_pool_indices, write_req_to_token=True)
        return self._commit(int(self.ret_values.item()), out_indices)

    def _commit(self, num_new_pages: int, out_indices: torch.Tensor) -> Optional[torch.Tensor]:
        if num_new_pages > len(self.free_pages):
            return None
        self.free_pages = self.free_pages[num_new_pages:]
        return out_indices
