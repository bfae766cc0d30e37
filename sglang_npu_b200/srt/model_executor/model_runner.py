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


@dataclass
class HostExtendBatch:
    """One chunked-prefill step as the scheduler hands it over (pinned host memory), schedule_batch.py:1180-1302:
    `input_ids` / `out_cache_loc` / `positions` hold the NEW tokens of every request back to back."""
    input_ids: torch.Tensor         # [T] int64
    req_pool_indices: torch.Tensor  # [bs] int64
    seq_lens: torch.Tensor          # [bs] int64 (prefix + new)
    prefix_lens: torch.Tensor       # [bs] int64 (radix-cache hit / earlier chunks)
    out_cache_loc: torch.Tensor     # [T] int64
    positions: torch.Tensor         # [T] int64

    def nbytes(self) -> int:
        return sum(t.numel() * t.element_size() for t in
                   (self.input_ids, self.req_pool_indices, self.seq_lens, self.prefix_lens, self.out_cache_loc,
                    self.positions))


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

    def forward_decode(self, input_ids, req_pool_indices, seq_lens, out_cache_loc, seq_lens_cpu=None,
                       shared_prefix_len: int = 0):
        fb = self._make_batch(ForwardMode.DECODE, input_ids, req_pool_indices, seq_lens, out_cache_loc, seq_lens_cpu,
                              shared_prefix_len=shared_prefix_len)
        fb.positions = seq_lens - 1  # forward_batch_info.py:960-962
        self.attn_backend.init_forward_metadata(fb)
        hidden = self.model(input_ids, fb.positions, fb)
        return self.model.logits(hidden)

    def forward_extend(self, input_ids, req_pool_indices, seq_lens, prefix_lens, out_cache_loc, positions,
                       seq_lens_cpu=None, extend_seq_lens_cpu=None, extend_prefix_lens_cpu=None, greedy=False):
        """`greedy=True` returns the sampled next-token id of every request instead of its last-token logits."""
        ext = (seq_lens - prefix_lens.to(seq_lens.dtype)).to(torch.int32)
        fb = self._make_batch(ForwardMode.EXTEND, input_ids, req_pool_indices, seq_lens, out_cache_loc, seq_lens_cpu,
                              extend_num_tokens=int(input_ids.numel()), extend_seq_lens=ext,
                              extend_prefix_lens=prefix_lens.to(torch.int32), extend_seq_lens_cpu=extend_seq_lens_cpu,
                              extend_prefix_lens_cpu=extend_prefix_lens_cpu)
        fb.positions = positions
        self.attn_backend.init_forward_metadata(fb)
        hidden = self.model(input_ids, positions, fb)
        last = torch.cumsum(ext, 0).long() - 1  # logits of the last token of every request
        if greedy:
            return self.model.greedy(hidden[last])
        return self.model.logits(hidden[last])

    def prefill_step(self, host_batch: "HostExtendBatch") -> torch.Tensor:
        """The public end-to-end call of one chunked-prefill step: pinned host index tensors in -> pinned host
        next-token ids out (synchronous).  Eager launches (the reference captures CUDA graphs for decode only)."""
        dev = self.device
        d = {k: getattr(host_batch, k).to(dev, non_blocking=True) for k in
             ("input_ids", "req_pool_indices", "seq_lens", "prefix_lens", "out_cache_loc", "positions")}
        ext = (host_batch.seq_lens - host_batch.prefix_lens).tolist()
        ids = self.forward_extend(d["input_ids"], d["req_pool_indices"], d["seq_lens"], d["prefix_lens"], d["out_cache_loc"],
                                  d["positions"], seq_lens_cpu=host_batch.seq_lens, extend_seq_lens_cpu=ext,
                                  extend_prefix_lens_cpu=host_batch.prefix_lens.tolist(), greedy=True)
        if getattr(self, "_host_prefill_out", None) is None or self._host_prefill_out.numel() != ids.numel():
            self._host_prefill_out = torch.zeros(ids.numel(), dtype=torch.int64).pin_memory()
        self._host_prefill_out.copy_(ids, non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()
        return self._host_prefill_out

    # ---- decode CUDA graph at a fixed batch size -------------------------------------------------------
    def capture_decode_graph(self, bs: int, shared_prefix_len: int = 0):
        """`shared_prefix_len` > 0: every request of the batches this graph will replay shares its first that many
        tokens (one radix-cache node): the shared-prefix decode is captured."""
        dev = self.device
        self.attn_backend.capture_shared_prefix_len = shared_prefix_len
        st = dict(
            input_ids=torch.zeros(bs, dtype=torch.int64, device=dev),
            req_pool_indices=torch.zeros(bs, dtype=torch.int64, device=dev),
            seq_lens=torch.full((bs,), self.attn_backend.get_cuda_graph_seq_len_fill_value(), dtype=torch.int64, device=dev),
            out_cache_loc=torch.zeros(bs, dtype=torch.int64, device=dev),
            positions=torch.zeros(bs, dtype=torch.int64, device=dev),
            next_token_ids=torch.zeros(bs, dtype=torch.int64, device=dev),
            keys=torch.zeros(bs, dtype=torch.int64, device=dev),
        )
        self.attn_backend.init_cuda_graph_state(bs, bs)
        fb = self._make_batch(ForwardMode.DECODE, st["input_ids"], st["req_pool_indices"], st["seq_lens"],
                              st["out_cache_loc"])
        fb.positions = st["positions"]

        def run_once():
            self.attn_backend.init_forward_metadata_capture_cuda_graph(bs, bs, st["req_pool_indices"], st["seq_lens"],
                                                                      None, ForwardMode.DECODE, None)
            # embedding gather + positions = seq_lens - 1 in one launch; lm_head on the weight-streaming GEMM; greedy
            # sampling (sampler.py: torch.argmax(logits, -1)) vocab-parallel at TP > 1 (8 bytes per row on the wire)
            hidden = self.model(st["input_ids"], st["positions"], fb, seq_lens_for_positions=st["seq_lens"])
            self.model.greedy(hidden, out=st["next_token_ids"], key_buf=st["keys"])

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

    # ---- a decode loop with the caller-side page-table kernels in it ---------------------------------------------------
    def attach_allocator(self, allocator) -> None:
        """`PagedTokenToKVPoolAllocator` whose free pages feed `decode_loop_step`."""
        self.allocator = allocator

    def decode_loop_step(self, host_batch: HostDecodeBatch) -> torch.Tensor:
        """One step of a real decode loop (schedule_batch.py:1519-1588 prepare_for_decode + the forward): the scheduler
        hands over pinned-host input_ids / req_pool_indices / seq_lens (already advanced to include the token being
        decoded); the KV slot of that token is allocated ON THE DEVICE by the fused get_last_loc + alloc_decode +
        write_req_to_token_pool launch (csrc/page_table.cu), then the captured graph runs and the sampled ids come back
        to pinned host memory.  host_batch.out_cache_loc is not read."""
        st = self._static
        st["input_ids"].copy_(host_batch.input_ids, non_blocking=True)
        st["req_pool_indices"].copy_(host_batch.req_pool_indices, non_blocking=True)
        st["seq_lens"].copy_(host_batch.seq_lens, non_blocking=True)
        loc = self.allocator.alloc_decode_fused(self.req_to_token_pool.req_to_token, st["req_pool_indices"], st["seq_lens"])
        if loc is None:
            raise RuntimeError("decode_loop_step: out of KV pages")
        st["out_cache_loc"].copy_(loc, non_blocking=True)
        self.graph.replay()
        self._host_out.copy_(st["next_token_ids"], non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        return self._host_out

    def decode_step(self, host_batch: HostDecodeBatch) -> torch.Tensor:
        """The public end-to-end call: pinned host indices in -> pinned host next-token ids out (synchronous)."""
        out = self.replay_decode(host_batch.input_ids, host_batch.req_pool_indices, host_batch.seq_lens,
                                 host_batch.out_cache_loc, host_batch.seq_lens)
        self._host_out.copy_(out, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        return self._host_out
