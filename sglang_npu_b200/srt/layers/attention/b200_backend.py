"""B200AttnBackend — the drop-in `AttentionBackend` for `--attention-backend b200`.

Boundary: reference python/sglang/srt/layers/attention/base_attn_backend.py:14-117 (interface),
triton_backend.py:41-122 / :160-336 / :338-627 / :632-732 (the backend this replaces on GPU),
torch_native_backend.py:182-267 (the oracle's semantics).

Differences from the Triton backend that matter to a maintainer:
  * no CSR `kv_indices` rebuild per step: the kernels read `req_to_token` rows directly, so
    `init_forward_metadata` for decode only picks the split-KV chunking (no kernel launch,
    no host sync) and CUDA-graph replay needs no metadata refresh at all;
  * `set_kv_buffer` is one fused scatter launch;
  * there is no Triton / FlashInfer / CPU fallback: tensors must live on a CUDA device.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Optional

import torch

from .... import kernels
from .base_attn_backend import AttentionBackend

_NUM_RESIDENT_CTAS_PER_SM = 2  # decode kernel: __launch_bounds__(128, 2), ~100 KiB smem per CTA


@dataclass
class ForwardMetadata:
    # decode
    kv_chunk: int = 0
    max_splits: int = 0
    workspace: Optional[torch.Tensor] = None
    shared_prefix_len: int = 0  # > 0: shared-prefix (cascade) decode, see forward_decode
    seg_len: int = 0
    # extend
    qo_indptr: Optional[torch.Tensor] = None
    prefix_lens: Optional[torch.Tensor] = None
    max_extend_len: int = 0
    # extend over an FP8 KV cache: staging of the up-cast prefix (kernels.fp8_kv_gather_dequant)
    stage_offsets: Optional[torch.Tensor] = None   # [bs] int32, row of prefix token 0 of request b
    stage_table: Optional[torch.Tensor] = None     # [bs, max_prefix] int32 page table into the staging pool
    stage_rows: int = 0
    max_prefix_len: int = 0


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


def plan_decode_splits(batch: int, num_kv_heads: int, group: int, max_len: int, sm_count: int,
                       forced_chunk: int = 0, max_chunk: int = 2048, exact: bool = True):
    """Pick (kv_chunk, max_splits) for the split-KV decode kernel.

    One CTA handles one (request, KV head, <=16 q heads, chunk); aim for ~3 waves of the
    2 x SM-count resident CTAs, with chunks of 256..2048 tokens (multiples of 16), and
    make the chunks of the longest request even.  Pure host arithmetic (graph-safe).
    Fewer, longer chunks win once the chip is covered a few times: every chunk costs a pipeline
    fill and an fp32 partial that stage 2 reads back (cfg 2, measured: 5 chunks of 832 tokens
    8.374 ms/step, 2 chunks of 2048 8.311 ms/step; kernel alone 0.1915 vs 0.1874 ms).
    exact=False: `max_len` is a worst-case bound, not the batch's longest request (graph capture / no host lengths) —
    evening the chunks against it would mis-size them for the requests that actually run (a 4k batch planned under a
    4k+32 bound got 3 x 1376 instead of 2 x 2048: +8 % kernel time).
    max_chunk: the FP8-KV kernel (3 CTAs/SM, half the bytes per token) is best at 1024 (0.118 vs 0.128 ms at 2048).
    """
    max_len = max(int(max_len), 1)
    if forced_chunk > 0:
        chunk = _round_up(forced_chunk, 16)
    else:
        units = max(batch * num_kv_heads * ((group + 15) // 16), 1)
        target = sm_count * _NUM_RESIDENT_CTAS_PER_SM * 3
        chunk = _round_up(max(max_len * units // target, 1), 64)
        if 4 * units >= 3 * sm_count * _NUM_RESIDENT_CTAS_PER_SM:
            # the (request, KV head) units alone already cover ~a wave: do not split further than max_chunk asks for
            # (bs 256 x 2048 tokens x 1 KV head: 1 chunk 0.064 ms, 4 chunks 0.071 ms, 8 chunks 0.075 ms)
            chunk = max_chunk
        chunk = min(max(chunk, 256), max_chunk)
        if exact:  # max_len is the longest request of THIS batch: make its chunks even
            splits = (max_len + chunk - 1) // chunk
            chunk = min(_round_up((max_len + splits - 1) // splits, 16), max_chunk)
        else:      # max_len is only a bound (CUDA-graph capture): keep whole tiles; splits past a request's end exit at once
            chunk = _round_up(chunk, 128)
    chunk = max(16, min(chunk, 4096))
    return chunk, (max_len + chunk - 1) // chunk


class B200AttnBackend(AttentionBackend):
    def __init__(self, model_runner, skip_prefill: bool = False):
        super().__init__()
        self.device = torch.device(model_runner.device)
        self._require_device(self.device)
        self.skip_prefill = skip_prefill
        self.req_to_token = model_runner.req_to_token_pool.req_to_token
        tp = getattr(model_runner, "tp_size", 1) or 1
        cfg = model_runner.model_config
        self.num_head = cfg.num_attention_heads // tp
        self.num_kv_head = cfg.get_num_kv_heads(tp)
        self.max_context_len = cfg.context_len
        self.sliding_window_size = getattr(model_runner, "sliding_window_size", None)
        self.head_dim = model_runner.token_to_kv_pool.get_key_buffer(0).shape[-1]
        self.v_head_dim = model_runner.token_to_kv_pool.get_value_buffer(0).shape[-1]
        if self.v_head_dim != self.head_dim:
            raise NotImplementedError("B200AttnBackend: v_head_dim != head_dim (MLA-style) is out of scope")
        self.sm_count = self._device_sm_count(self.device)
        self.forced_chunk = int(os.environ.get("B200_DECODE_KV_CHUNK", "0"))
        self.min_shared_prefix = int(os.environ.get("B200_MIN_SHARED_PREFIX", "1024"))
        self._max_chunk = 1024 if model_runner.token_to_kv_pool.dtype == torch.float8_e4m3fn else 2048
        self.forward_metadata: Optional[ForwardMetadata] = None
        self.capture_shared_prefix_len = 0
        self._kv_dtype = model_runner.token_to_kv_pool.dtype
        self._workspace: Optional[torch.Tensor] = None
        self._graph_workspace: Optional[torch.Tensor] = None
        self._graph_max_bs = 0
        self._stage_kv = None
        self._stage_rows_idx = None
        self._sp_stage_meta = None
        self._sp_stage_kv = None
        max_bs = model_runner.req_to_token_pool.size
        self.qo_indptr = torch.zeros((max_bs + 1,), dtype=torch.int32, device=self.device)

    @staticmethod
    def _require_device(device: torch.device) -> None:
        if device.type != "cuda":
            raise RuntimeError("B200AttnBackend requires a CUDA device (no CPU fallback in this backend)")
        kernels._cabi.load()  # fail loudly at construction if the extension is missing

    @staticmethod
    def _device_sm_count(device: torch.device) -> int:
        return torch.cuda.get_device_properties(device).multi_processor_count

    # ---- metadata -------------------------------------------------------------------------
    def _shared_prefix_supported(self) -> bool:
        return self._kv_dtype in (torch.bfloat16, torch.float16, torch.float8_e4m3fn)

    def _shared_prefix_stage(self, shared: int, index_dtype: torch.dtype):
        """FP8 KV cache: the index tensors of the staged shared prefix (device-side fills only: capturable)."""
        st = self._sp_stage_meta
        if st is None or st["shared"] != shared or st["index"].dtype != index_dtype:
            st = dict(shared=shared,
                      prefix_lens=torch.full((1,), shared, dtype=torch.int32, device=self.device),
                      offsets=torch.zeros((1,), dtype=torch.int32, device=self.device),
                      table=torch.arange(shared, dtype=torch.int32, device=self.device),
                      index=torch.zeros((1,), dtype=index_dtype, device=self.device))
            self._sp_stage_meta = st
        return st

    def _decode_workspace(self, bs: int, max_splits: int, graph: bool = False) -> Optional[torch.Tensor]:
        if max_splits <= 1:
            return None
        need = kernels.decode_attn_workspace_bytes(bs, self.num_head, self.head_dim, max_splits)
        cur = self._graph_workspace if graph else self._workspace
        if cur is None or cur.numel() < need:
            cur = torch.empty(need, dtype=torch.uint8, device=self.device)
            if graph:
                self._graph_workspace = cur
            else:
                self._workspace = cur
        return cur

    def init_forward_metadata(self, forward_batch):
        mode = forward_batch.forward_mode
        bs = forward_batch.batch_size
        group = self.num_head // self.num_kv_head
        if mode.is_decode_or_idle():
            exact = forward_batch.seq_lens_cpu is not None
            if exact:
                max_len = int(forward_batch.seq_lens_cpu.max()) if bs > 0 else 1
            else:
                max_len = self.max_context_len  # no host sync: plan for the worst case
            # Extension over the reference's ForwardBatch: a scheduler whose radix cache matched the same prefix for
            # every request of the batch (BASELINE config 5) sets `shared_prefix_len`; the first that many entries of
            # every req_to_token row are then the same physical slots.  Worth the extra launch from ~1k tokens on.
            shared = int(getattr(forward_batch, "shared_prefix_len", 0) or 0)
            if shared and forward_batch.seq_lens_cpu is not None and bs > 0:
                shared = min(shared, int(forward_batch.seq_lens_cpu.min()) - 1)  # every request keeps >= 1 own token
            if shared >= self.min_shared_prefix and bs >= 2 and self._shared_prefix_supported():
                shared = shared // 128 * 128  # whole KV tiles for the prefix pass; the remainder stays per-request
                chunk, splits = plan_decode_splits(bs, self.num_kv_head, group, max(max_len - shared, 1),
                                                   self.sm_count, self.forced_chunk, self._max_chunk, exact)
                seg = kernels.plan_shared_prefix_segments(shared, bs, self.num_head, self.sm_count, head_dim=self.head_dim)
                n_seg = (shared + seg - 1) // seg
                self.forward_metadata = ForwardMetadata(
                    kv_chunk=chunk, max_splits=splits, shared_prefix_len=shared, seg_len=seg,
                    workspace=self._decode_workspace(bs, n_seg + splits))
                return
            chunk, splits = plan_decode_splits(bs, self.num_kv_head, group, max_len, self.sm_count,
                                               self.forced_chunk, self._max_chunk, exact)
            self.forward_metadata = ForwardMetadata(kv_chunk=chunk, max_splits=splits,
                                                    workspace=self._decode_workspace(bs, splits))
        else:
            ext = forward_batch.extend_seq_lens
            qo_indptr = self.qo_indptr[: bs + 1]
            qo_indptr[1:] = torch.cumsum(ext, dim=0)
            if forward_batch.extend_seq_lens_cpu is not None:
                max_extend = max(forward_batch.extend_seq_lens_cpu) if bs > 0 else 0
            else:
                max_extend = int(ext.max().item())  # same host sync as triton_backend.py:325
            prefix_lens = forward_batch.extend_prefix_lens.to(torch.int32)
            self.forward_metadata = ForwardMetadata(qo_indptr=qo_indptr, prefix_lens=prefix_lens,
                                                    max_extend_len=int(max_extend))
            if forward_batch.token_to_kv_pool.dtype == torch.float8_e4m3fn:
                md = self.forward_metadata
                if forward_batch.extend_prefix_lens_cpu is not None:
                    md.max_prefix_len = max(forward_batch.extend_prefix_lens_cpu) if bs > 0 else 0
                    md.stage_rows = sum(forward_batch.extend_prefix_lens_cpu)
                else:
                    md.max_prefix_len = int(prefix_lens.max().item())
                    md.stage_rows = int(prefix_lens.sum().item())
                incl = torch.cumsum(prefix_lens, dim=0, dtype=torch.int32)
                md.stage_offsets = incl - prefix_lens
                md.stage_table = (md.stage_offsets.view(-1, 1)
                                  + torch.arange(max(md.max_prefix_len, 1), dtype=torch.int32, device=self.device).view(1, -1))
                self._stage_rows_idx = torch.arange(bs, dtype=forward_batch.req_pool_indices.dtype, device=self.device)

    # ---- CUDA graphs (base_attn_backend.py:22-55; cuda_graph_runner.py:283-288, :642, :773) --------
    def init_cuda_graph_state(self, max_bs: int, max_num_tokens: int):
        self._graph_max_bs = max_bs
        group = self.num_head // self.num_kv_head
        _, splits = plan_decode_splits(1, self.num_kv_head, group, self.max_context_len, self.sm_count,
                                       self.forced_chunk, self._max_chunk, exact=False)
        self._decode_workspace(max_bs, max(splits, 2), graph=True)

    def init_forward_metadata_capture_cuda_graph(self, bs: int, num_tokens: int, req_pool_indices, seq_lens,
                                                 encoder_lens, forward_mode, spec_info):
        if not forward_mode.is_decode_or_idle():
            raise NotImplementedError("B200AttnBackend: CUDA graphs are captured for DECODE only")
        group = self.num_head // self.num_kv_head
        # Extension: `capture_shared_prefix_len` (set by the graph runner before capture) captures the shared-prefix
        # decode for a batch whose radix-matched common prefix has that many tokens; 16-bit and e4m3 KV caches alike.
        shared = int(self.capture_shared_prefix_len or 0) // 128 * 128
        if shared >= self.min_shared_prefix and bs >= 2 and self._shared_prefix_supported():
            chunk, splits = plan_decode_splits(bs, self.num_kv_head, group, max(self.max_context_len - shared, 1),
                                               self.sm_count, self.forced_chunk, self._max_chunk, exact=False)
            seg = kernels.plan_shared_prefix_segments(shared, bs, self.num_head, self.sm_count, head_dim=self.head_dim)
            n_seg = (shared + seg - 1) // seg
            self.forward_metadata = ForwardMetadata(kv_chunk=chunk, max_splits=splits, shared_prefix_len=shared, seg_len=seg,
                                                    workspace=self._decode_workspace(bs, n_seg + splits, graph=True))
            return
        chunk, splits = plan_decode_splits(bs, self.num_kv_head, group, self.max_context_len, self.sm_count,
                                           self.forced_chunk, self._max_chunk, exact=False)
        self.forward_metadata = ForwardMetadata(kv_chunk=chunk, max_splits=splits,
                                                workspace=self._decode_workspace(bs, splits, graph=True))

    def init_forward_metadata_replay_cuda_graph(self, bs: int, req_pool_indices, seq_lens, seq_lens_sum: int,
                                                encoder_lens, forward_mode, spec_info, seq_lens_cpu):
        # Nothing to refresh: the captured kernels read req_pool_indices / seq_lens / req_to_token
        # from the graph runner's static tensors, and the chunking was fixed at capture.
        return

    def get_cuda_graph_seq_len_fill_value(self):
        return 1

    # ---- per-layer entry points -----------------------------------------------------------------
    def _save_kv(self, layer, forward_batch, k, v):
        pool = forward_batch.token_to_kv_pool
        kb = pool.get_key_buffer(layer.layer_id)
        if k.dtype == kb.dtype and layer.k_scale is None:
            kernels.set_kv_buffer(kb, pool.get_value_buffer(layer.layer_id), forward_batch.out_cache_loc, k, v)
        else:  # dtype conversion / scaling rules live in the pool (memory_pool.py:385-391)
            pool.set_kv_buffer(layer, forward_batch.out_cache_loc, k, v, *self._kv_scales(layer))

    @staticmethod
    def _kv_scales(layer):
        """(k_scale, v_scale) as Python floats or None (radix_attention.py:73-76 keeps both a tensor and a float)."""
        ks = getattr(layer, "k_scale_float", None)
        vs = getattr(layer, "v_scale_float", None)
        if ks is None and layer.k_scale is not None:
            ks = float(layer.k_scale)
        if vs is None and layer.v_scale is not None:
            vs = float(layer.v_scale)
        return ks, vs

    def _check_layer(self, layer):
        if layer.sliding_window_size is not None and layer.sliding_window_size > -1:
            raise NotImplementedError("B200AttnBackend: sliding-window attention is out of scope")
        if layer.qk_head_dim != layer.v_head_dim:
            raise NotImplementedError("B200AttnBackend: qk_head_dim != v_head_dim is out of scope")

    def forward_decode(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True, int8_out: bool = False,
                       kv_in_cache: bool = False):
        """int8_out=True returns the output already per-token int8-quantised (kernels.Int8Activation) for a W8A8-int8
        o_proj: the quant is fused into the split-KV merge kernel."""
        self._check_layer(layer)
        q = q.reshape(-1, layer.tp_q_head_num * layer.qk_head_dim)
        o = torch.empty_like(q)
        if save_kv_cache:
            self._save_kv(layer, forward_batch, k, v)
        md = self.forward_metadata
        pool = forward_batch.token_to_kv_pool
        if md.shared_prefix_len > 0:
            extra = {}
            if pool.dtype != q.dtype:
                # FP8 KV cache: up-cast the shared tokens once per layer into a 16-bit staging copy (2 x prefix x Hkv x D
                # bytes read, twice that written: ~4 us at config 5 per GPU) for the tensor-core prefix pass; every
                # request's own tokens are read as e4m3 by the FP8-KV decode kernel
                if pool.dtype != torch.float8_e4m3fn:
                    raise NotImplementedError(f"B200AttnBackend: KV cache dtype {pool.dtype} is not supported")
                shared = md.shared_prefix_len
                st = self._shared_prefix_stage(shared, forward_batch.req_pool_indices.dtype)
                shape = (shared, layer.tp_k_head_num, layer.qk_head_dim)
                if self._sp_stage_kv is None or self._sp_stage_kv[0].shape != shape or self._sp_stage_kv[0].dtype != q.dtype:
                    self._sp_stage_kv = (torch.empty(shape, dtype=q.dtype, device=q.device),
                                         torch.empty(shape, dtype=q.dtype, device=q.device))
                sk, sv = self._sp_stage_kv
                ks, vs = self._kv_scales(layer)
                kernels.fp8_kv_gather_dequant(sk, sv, pool.get_key_buffer(layer.layer_id),
                                              pool.get_value_buffer(layer.layer_id),
                                              forward_batch.req_to_token_pool.req_to_token,
                                              forward_batch.req_pool_indices[:1], st["prefix_lens"], st["offsets"], shared,
                                              ks, vs)
                extra = dict(stage=(sk, sv, st["table"], st["index"]), k_scale=ks, v_scale=vs)
            act = kernels.decode_attention_shared_prefix_fwd(
                q.view(-1, layer.tp_q_head_num, layer.qk_head_dim),
                pool.get_key_buffer(layer.layer_id), pool.get_value_buffer(layer.layer_id),
                o.view(-1, layer.tp_q_head_num, layer.v_head_dim),
                forward_batch.req_to_token_pool.req_to_token, forward_batch.req_pool_indices,
                forward_batch.seq_lens, md.shared_prefix_len, md.seg_len, md.workspace, md.kv_chunk, md.max_splits,
                layer.scaling, layer.logit_cap or 0.0, int8_out=int8_out, **extra)
            return act if int8_out else o
        act = kernels.decode_attention_fwd(
            q.view(-1, layer.tp_q_head_num, layer.qk_head_dim),
            pool.get_key_buffer(layer.layer_id), pool.get_value_buffer(layer.layer_id),
            o.view(-1, layer.tp_q_head_num, layer.v_head_dim),
            forward_batch.req_to_token_pool.req_to_token, forward_batch.req_pool_indices,
            forward_batch.seq_lens, md.workspace, md.kv_chunk, md.max_splits,
            layer.scaling, layer.logit_cap or 0.0, int8_out=int8_out,
            **({} if pool.dtype == q.dtype else dict(zip(("k_scale", "v_scale"), self._kv_scales(layer)))))
        return act if int8_out else o

    def forward_extend(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True, int8_out: bool = False,
                       kv_in_cache: bool = False):
        """kv_in_cache=True (with save_kv_cache=False): the caller has ALREADY written this chunk's K/V into the pool
        (the fused rope + KV-write launch), so the kernel reads new and cached tokens alike through the page table — the
        same path as after its own write — instead of gathering the new tokens from the strided qkv activations."""
        self._check_layer(layer)
        q = q.reshape(-1, layer.tp_q_head_num * layer.qk_head_dim)
        o = torch.empty_like(q)
        if save_kv_cache:
            self._save_kv(layer, forward_batch, k, v)
        md = self.forward_metadata
        pool = forward_batch.token_to_kv_pool
        causal = not (layer.is_cross_attention or getattr(layer.attn_type, "value", "") == "encoder_only")
        if pool.dtype != q.dtype:
            # FP8 KV cache: the cached prefix is up-cast into a 16-bit staging pool (one launch, prefill is
            # tensor-bound so its traffic is noise) and the new tokens are read from the activations un-quantised —
            # what the Triton extend kernel does with an fp8 k_buffer and bf16 k_extend (extend_attention.py:302-440)
            if pool.dtype != torch.float8_e4m3fn:
                raise NotImplementedError(f"B200AttnBackend: KV cache dtype {pool.dtype} is not supported")
            rows = max(md.stage_rows, 1)
            if self._stage_kv is None or self._stage_kv[0].shape[0] < rows or self._stage_kv[0].dtype != q.dtype:
                shape = (rows, layer.tp_k_head_num, layer.qk_head_dim)
                self._stage_kv = (torch.empty(shape, dtype=q.dtype, device=q.device),
                                  torch.empty(shape, dtype=q.dtype, device=q.device))
            sk, sv = self._stage_kv
            ks, vs = self._kv_scales(layer)
            kernels.fp8_kv_gather_dequant(sk, sv, pool.get_key_buffer(layer.layer_id),
                                          pool.get_value_buffer(layer.layer_id),
                                          forward_batch.req_to_token_pool.req_to_token, forward_batch.req_pool_indices,
                                          md.prefix_lens, md.stage_offsets, md.max_prefix_len, ks, vs)
            kernels.extend_attention_fwd(
                q.view(-1, layer.tp_q_head_num, layer.qk_head_dim),
                k.view(-1, layer.tp_k_head_num, layer.qk_head_dim), v.view(-1, layer.tp_v_head_num, layer.v_head_dim),
                o.view(-1, layer.tp_q_head_num, layer.v_head_dim), sk, sv, md.stage_table, self._stage_rows_idx,
                md.qo_indptr, md.prefix_lens, md.max_extend_len, layer.scaling, layer.logit_cap or 0.0, causal)
            if int8_out:
                xq, xs = kernels.per_token_quant_int8(o)
                return kernels.Int8Activation(xq, xs, o.dtype)
            return o
        kernels.extend_attention_fwd(
            q.view(-1, layer.tp_q_head_num, layer.qk_head_dim),
            None if (save_kv_cache or kv_in_cache) else k.view(-1, layer.tp_k_head_num, layer.qk_head_dim),
            None if (save_kv_cache or kv_in_cache) else v.view(-1, layer.tp_v_head_num, layer.v_head_dim),
            o.view(-1, layer.tp_q_head_num, layer.v_head_dim),
            pool.get_key_buffer(layer.layer_id), pool.get_value_buffer(layer.layer_id),
            forward_batch.req_to_token_pool.req_to_token, forward_batch.req_pool_indices,
            md.qo_indptr, md.prefix_lens, md.max_extend_len,
            layer.scaling, layer.logit_cap or 0.0, causal)
        if int8_out:  # prefill is tensor-bound; the quant stays its own (bandwidth-bound) kernel there
            xq, xs = kernels.per_token_quant_int8(o)
            return kernels.Int8Activation(xq, xs, o.dtype)
        return o

    def support_triton(self):
        return False
