"""ForwardMode / ForwardBatch — the fields of the reference bundle that the hot path reads
(reference python/sglang/srt/model_executor/forward_batch_info.py:68-143, :164-300).

The B200 attention backend is duck-typed: it works with this dataclass or with the
reference's own ForwardBatch (same field names).
"""
from __future__ import annotations

from dataclasses import dataclass
from enum import IntEnum, auto
from typing import Any, List, Optional

import torch


class ForwardMode(IntEnum):
    EXTEND = auto()
    DECODE = auto()
    MIXED = auto()
    IDLE = auto()
    TARGET_VERIFY = auto()
    DRAFT_EXTEND = auto()
    DUMMY_FIRST = auto()
    SPLIT_PREFILL = auto()

    def is_prefill(self):
        return self.is_extend()

    def is_extend(self):
        return self in (ForwardMode.EXTEND, ForwardMode.MIXED, ForwardMode.DRAFT_EXTEND,
                        ForwardMode.TARGET_VERIFY)

    def is_decode(self):
        return self == ForwardMode.DECODE

    def is_mixed(self):
        return self == ForwardMode.MIXED

    def is_idle(self):
        return self == ForwardMode.IDLE

    def is_decode_or_idle(self):
        return self in (ForwardMode.DECODE, ForwardMode.IDLE)

    def is_target_verify(self):
        return self == ForwardMode.TARGET_VERIFY

    def is_draft_extend(self):
        return self == ForwardMode.DRAFT_EXTEND

    def is_cuda_graph(self):
        return self in (ForwardMode.DECODE, ForwardMode.TARGET_VERIFY, ForwardMode.IDLE)


@dataclass
class ForwardBatch:
    """Inputs of one forward pass (subset of the reference fields, same names)."""

    forward_mode: ForwardMode
    batch_size: int
    input_ids: Optional[torch.Tensor]
    req_pool_indices: torch.Tensor   # [bs] int64 (int32 accepted)
    seq_lens: torch.Tensor           # [bs] int64, INCLUDING the token(s) being processed
    out_cache_loc: torch.Tensor      # [T] int64 KV slots of the tokens being processed
    seq_lens_sum: int
    seq_lens_cpu: Optional[torch.Tensor] = None
    positions: Optional[torch.Tensor] = None

    # extend only
    extend_num_tokens: Optional[int] = None
    extend_seq_lens: Optional[torch.Tensor] = None     # [bs] int32
    extend_prefix_lens: Optional[torch.Tensor] = None  # [bs] int32
    extend_start_loc: Optional[torch.Tensor] = None
    extend_prefix_lens_cpu: Optional[List[int]] = None
    extend_seq_lens_cpu: Optional[List[int]] = None

    # pools + backend
    req_to_token_pool: Any = None
    token_to_kv_pool: Any = None
    attn_backend: Any = None

    spec_info: Any = None
    encoder_lens: Optional[torch.Tensor] = None

    # Extension (not a reference field): number of leading tokens whose KV slots are identical for EVERY request of a
    # decode batch (a radix-cache hit on a common prefix).  B200AttnBackend then runs the shared-prefix decode.
    shared_prefix_len: int = 0
