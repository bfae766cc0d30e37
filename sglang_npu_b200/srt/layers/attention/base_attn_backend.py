"""AttentionBackend plugin interface — same methods, arguments and dispatch rule as the
reference base class (python/sglang/srt/layers/attention/base_attn_backend.py:14-117), so a
subclass written against this file drops into the reference unchanged.
"""
from __future__ import annotations

from abc import ABC, abstractmethod
from typing import Optional

import torch


class AttentionBackend(ABC):
    @abstractmethod
    def init_forward_metadata(self, forward_batch):
        """Prepare per-forward metadata (called once per forward, model_runner.py:1553-1573)."""
        raise NotImplementedError()

    def init_cuda_graph_state(self, max_bs: int, max_num_tokens: int):
        raise NotImplementedError()

    def init_forward_metadata_capture_cuda_graph(self, bs: int, num_tokens: int,
                                                 req_pool_indices: torch.Tensor, seq_lens: torch.Tensor,
                                                 encoder_lens: Optional[torch.Tensor], forward_mode,
                                                 spec_info):
        raise NotImplementedError()

    def init_forward_metadata_replay_cuda_graph(self, bs: int, req_pool_indices: torch.Tensor,
                                                seq_lens: torch.Tensor, seq_lens_sum: int,
                                                encoder_lens: Optional[torch.Tensor], forward_mode,
                                                spec_info, seq_lens_cpu: Optional[torch.Tensor]):
        raise NotImplementedError()

    def get_cuda_graph_seq_len_fill_value(self):
        raise NotImplementedError()

    def forward(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True, **kwargs):
        """Mode dispatch of base_attn_backend.py:57-89."""
        if forward_batch.forward_mode.is_idle():
            return q.new_empty(q.shape[0], layer.tp_q_head_num * layer.v_head_dim)
        elif forward_batch.forward_mode.is_decode():
            return self.forward_decode(q, k, v, layer, forward_batch, save_kv_cache=save_kv_cache, **kwargs)
        else:
            return self.forward_extend(q, k, v, layer, forward_batch, save_kv_cache=save_kv_cache, **kwargs)

    def forward_decode(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True):
        raise NotImplementedError()

    def forward_extend(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True):
        raise NotImplementedError()

    def support_triton(self):
        return True
