"""HotPathRunner — the minimal ModelRunner + CudaGraphRunner around the hot path.

Follows the reference's control flow for one forward (python/sglang/srt/model_executor/model_runner.py:
1547-1573 forward_decode / forward_extend -> attn_backend.init_forward_metadata -> model.forward ->
sample :1717) and its decode CUDA-graph protocol (cuda_graph_runner.py:283-288 init_cuda_graph_state,
:642 capture metadata, :773-792 replay), for a single fixed batch size.  The public call a user makes
is `decode_step(host_batch)`: pinned host index tensors in, pinned host next-token ids out.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import torch

from ..layers.attention.b200_backend import B200AttnBackend
from ..mem_cache.memory_pool import MHATokenToKVPool, ReqToTokenPool
from ... import kernels
from .forward_batch_info import ForwardBatch, ForwardMode
from .layer_stack import LayerStack, StackConfig, load_synthetic_weights


@dataclass
class HostDecodeBatch:
    """What the scheduler hands over per decode step (pinned host memory), schedule_batch.py:1519-1588."""
    input_ids: torch.Tensor         # [bs] int64
    req_pool_indices: torch.Tensor  # [bs] int64
    seq_lens: torch.Tensor          # [bs] int64 (including the token being decoded)
    out_cache_loc: torch.Tensor     # [bs] int64

    def nbytes(self) -> int:
        return sum(t.numel() * t.element_size() for t in
                   (self.input_ids, self.req_pool_indices, self.seq_lens, self.out_cache_loc))


class HotPathRunner:
    def __init__(self, cfg: StackConfig, quant_config, *, max_reqs: int, max_total_tokens: int, page_size: int,
                 context_len: Optional[int] = None, dtype=torch.bfloat16, device="cuda", num_layers: Optional[int] = None,
                 tp_size: int = 1, seed: int = 1234, weight_gain: float = 1.0, tp_comm_tokens: int = 0,
                 kv_cache_dtype: Optional[torch.dtype] = None):
        self.device = torch.device(device)
        self.model_config = cfg
        if context_len is not None:
            cfg.context_len = context_len
        self.tp_size = tp_size
        self.page_size = page_size
        self.dtype = dtype
        self.sliding_window_size = None
        self.server_args = None
        self.gpu_id = self.device.index or 0
        n_layers = num_layers if num_layers is not None else cfg.num_layers
        self.req_to_token_pool = ReqToTokenPool(max_reqs, cfg.context_len + 4, str(self.device))
        self.token_to_kv_pool = MHATokenToKVPool(max_total_tokens, page_size, kv_cache_dtype or dtype,  # --kv-cache-dtype
                                                 cfg.get_num_kv_heads(tp_size),
                                                 cfg.head_dim, n_layers, str(self.device))
        with torch.device(self.device):
            self.model = LayerStack(cfg, quant_config, dtype, n_layers)
        load_synthetic_weights(self.model, seed=seed, device=self.device, gain=weight_gain)
        self.attn_backend = B200AttnBackend(self)
        if tp_size > 1:
            from ..distributed.parallel_state import init_tp_comm

            self.tp_comm = init_tp_comm(max(max_reqs, tp_comm_tokens), cfg.hidden_size, dtype, self.device)
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self._static = None

    # ---- eager paths -------------------------------------------------------------------------------
    def _make_batch(self, mode, input_ids, req_pool_indices, seq_lens, out_cache_loc, seq_lens_cpu=None, **extend):
        return ForwardBatch(forward_mode=mode, batch_size=req_pool_indices.numel(), input_ids=input_ids,
                            req_pool_indices=req_pool_indices, seq_lens=seq_lens, out_cache_loc=out_cache_loc,
                            seq_lens_sum=0, seq_lens_cpu=seq_lens_cpu, req_to_token_pool=self.req_to_token_pool,
                            token_to_kv_pool=self.token_to_kv_pool, attn_backend=self.attn_backend, **extend)

    def forward_decode(self, input_ids, req_pool_indices, seq_lens, out_cache_loc, seq_lens_cpu=None):
        fb = self._make_batch(ForwardMode.DECODE, input_ids, req_pool_indices, seq_lens, out_cache_loc, seq_lens_cpu)
        fb.positions = seq_lens - 1  # forward_batch_info.py:960-962
        self.attn_backend.init_forward_metadata(fb)
        hidden = self.model(input_ids, fb.positions, fb)
        return self.model.logits(hidden)

    def forward_extend(self, input_ids, req_pool_indices, seq_lens, prefix_lens, out_cache_loc, positions,
                       seq_lens_cpu=None, extend_seq_lens_cpu=None):
        ext = (seq_lens - prefix_lens.to(seq_lens.dtype)).to(torch.int32)
        fb = self._make_batch(ForwardMode.EXTEND, input_ids, req_pool_indices, seq_lens, out_cache_loc, seq_lens_cpu,
                              extend_num_tokens=int(input_ids.numel()), extend_seq_lens=ext,
                              extend_prefix_lens=prefix_lens.to(torch.int32), extend_seq_lens_cpu=extend_seq_lens_cpu)
        fb.positions = positions
        self.attn_backend.init_forward_metadata(fb)
        hidden = self.model(input_ids, positions, fb)
        last = torch.cumsum(ext, 0).long() - 1  # logits of the last token of every request
        return self.model.logits(hidden[last])

    # ---- decode CUDA graph at a fixed batch size -------------------------------------------------------
    def capture_decode_graph(self, bs: int):
        dev = self.device
        st = dict(
            input_ids=torch.zeros(bs, dtype=torch.int64, device=dev),
            req_pool_indices=torch.zeros(bs, dtype=torch.int64, device=dev),
            seq_lens=torch.full((bs,), self.attn_backend.get_cuda_graph_seq_len_fill_value(), dtype=torch.int64, device=dev),
            out_cache_loc=torch.zeros(bs, dtype=torch.int64, device=dev),
            positions=torch.zeros(bs, dtype=torch.int64, device=dev),
            next_token_ids=torch.zeros(bs, dtype=torch.int64, device=dev),
        )
        self.attn_backend.init_cuda_graph_state(bs, bs)
        fb = self._make_batch(ForwardMode.DECODE, st["input_ids"], st["req_pool_indices"], st["seq_lens"],
                              st["out_cache_loc"])
        fb.positions = st["positions"]

        def run_once():
            self.attn_backend.init_forward_metadata_capture_cuda_graph(bs, bs, st["req_pool_indices"], st["seq_lens"],
                                                                      None, ForwardMode.DECODE, None)
            torch.sub(st["seq_lens"], 1, out=st["positions"])
            hidden = self.model(st["input_ids"], st["positions"], fb)
            logits = self.model.logits(hidden)
            kernels.argmax(logits, out=st["next_token_ids"])  # greedy sampling (sampler.py: torch.argmax(logits, -1))

        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(2):  # warm up allocator / lazy inits outside capture
                run_once()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            run_once()
        self.graph, self._static = graph, st
        self._host_out = torch.zeros(bs, dtype=torch.int64).pin_memory()
        return graph

    def replay_decode(self, input_ids, req_pool_indices, seq_lens, out_cache_loc, seq_lens_cpu=None):
        """Device-resident inputs -> device next-token ids (cuda_graph_runner.py:773-792)."""
        st = self._static
        st["input_ids"].copy_(input_ids, non_blocking=True)
        st["req_pool_indices"].copy_(req_pool_indices, non_blocking=True)
        st["seq_lens"].copy_(seq_lens, non_blocking=True)
        st["out_cache_loc"].copy_(out_cache_loc, non_blocking=True)
        self.attn_backend.init_forward_metadata_replay_cuda_graph(st["seq_lens"].numel(), st["req_pool_indices"],
                                                                  st["seq_lens"], 0, None, ForwardMode.DECODE, None,
                                                                  seq_lens_cpu)
        self.graph.replay()
        return st["next_token_ids"]

    def decode_step(self, host_batch: HostDecodeBatch) -> torch.Tensor:
        """The public end-to-end call: pinned host indices in -> pinned host next-token ids out (synchronous)."""
        out = self.replay_decode(host_batch.input_ids, host_batch.req_pool_indices, host_batch.seq_lens,
                                 host_batch.out_cache_loc, host_batch.seq_lens)
        self._host_out.copy_(out, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        return self._host_out
