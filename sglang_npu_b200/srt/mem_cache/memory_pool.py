"""Request->token page table and the paged MHA KV pool — the HBM-resident state the hot path
reads and writes (reference python/sglang/srt/mem_cache/memory_pool.py:47-94, :162-417).

Layout is identical to the reference so its scheduler/allocators can drive these pools:
  req_to_token : [max_num_reqs, max_context_len] int32, zero-initialised (stale -> slot 0)
  k_buffer[l], v_buffer[l] : [size + page_size, Hkv_local, D] in kv dtype; slots
  0..page_size-1 are the padding page that padded tokens write to.
`set_kv_buffer` is one fused scatter launch (csrc/kv_write.cu) instead of two index_put_.
"""
from __future__ import annotations

from typing import List, Optional, Union

import torch

from ... import kernels


class ReqToTokenPool:
    def __init__(self, size: int, max_context_len: int, device: str, enable_memory_saver: bool = False):
        self.size = size
        self.max_context_len = max_context_len
        self.device = device
        self.req_to_token = torch.zeros((size, max_context_len), dtype=torch.int32, device=device)
        self.free_slots = list(range(size))

    def write(self, indices, values):
        self.req_to_token[indices] = values

    def available_size(self):
        return len(self.free_slots)

    def alloc(self, need_size: int) -> Optional[List[int]]:
        if need_size > len(self.free_slots):
            return None
        select_index = self.free_slots[:need_size]
        self.free_slots = self.free_slots[need_size:]
        return select_index

    def free(self, free_index: Union[int, List[int]]):
        if isinstance(free_index, int):
            self.free_slots.append(free_index)
        else:
            self.free_slots.extend(free_index)

    def clear(self):
        self.free_slots = list(range(self.size))


class MHATokenToKVPool:
    def __init__(self, size: int, page_size: int, dtype: torch.dtype, head_num: int, head_dim: int,
                 layer_num: int, device: str, enable_memory_saver: bool = False,
                 start_layer: Optional[int] = None, end_layer: Optional[int] = None):
        self.size = size
        self.page_size = page_size
        self.dtype = dtype
        self.device = device
        # fp8 KV is stored as uint8 because index_put is not implemented for fp8 (memory_pool.py:114-118)
        self.store_dtype = torch.uint8 if dtype in (torch.float8_e5m2, torch.float8_e4m3fn) else dtype
        self.layer_num = layer_num
        self.start_layer = start_layer or 0
        self.end_layer = end_layer or layer_num - 1
        self.head_num = head_num
        self.head_dim = head_dim
        self.layer_transfer_counter = None
        self._create_buffers()

    def _create_buffers(self):
        shape = (self.size + self.page_size, self.head_num, self.head_dim)
        self.k_buffer = [torch.zeros(shape, dtype=self.store_dtype, device=self.device) for _ in range(self.layer_num)]
        self.v_buffer = [torch.zeros(shape, dtype=self.store_dtype, device=self.device) for _ in range(self.layer_num)]

    def get_kv_size_bytes(self):
        k = sum(x.numel() * x.element_size() for x in self.k_buffer)
        v = sum(x.numel() * x.element_size() for x in self.v_buffer)
        return k, v

    def get_key_buffer(self, layer_id: int):
        buf = self.k_buffer[layer_id - self.start_layer]
        return buf.view(self.dtype) if self.store_dtype != self.dtype else buf

    def get_value_buffer(self, layer_id: int):
        buf = self.v_buffer[layer_id - self.start_layer]
        return buf.view(self.dtype) if self.store_dtype != self.dtype else buf

    def get_kv_buffer(self, layer_id: int):
        return self.get_key_buffer(layer_id), self.get_value_buffer(layer_id)

    def set_kv_buffer(self, layer, loc: torch.Tensor, cache_k: torch.Tensor, cache_v: torch.Tensor,
                      k_scale: Optional[float] = None, v_scale: Optional[float] = None,
                      layer_id_override: Optional[int] = None):
        layer_id = layer_id_override if layer_id_override is not None else layer.layer_id
        kb = self.k_buffer[layer_id - self.start_layer]
        vb = self.v_buffer[layer_id - self.start_layer]
        if cache_k.dtype != self.dtype:
            # the cast branch (memory_pool.py:385-391) runs inside the scatter kernel; the reference's in-place
            # `cache_k.div_(k_scale)` side effect on the caller's tensor is not reproduced
            if self.dtype != torch.float8_e4m3fn:
                raise NotImplementedError(f"B200 KV pool: cast {cache_k.dtype} -> {self.dtype} is not supported "
                                          "(bf16/f16 -> float8_e4m3fn only)")
            kernels.set_kv_buffer(kb, vb, loc, cache_k, cache_v, k_scale, v_scale)
            return
        if self.store_dtype != self.dtype:
            cache_k = cache_k.view(self.store_dtype)
            cache_v = cache_v.view(self.store_dtype)
        kernels.set_kv_buffer(kb, vb, loc, cache_k, cache_v)
