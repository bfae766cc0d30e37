"""The caller loop around the hot path: a Llama / Qwen2 decoder stack built ONLY from the hot-path
pieces (parallel linears + quant methods, RadixAttention + B200AttnBackend, fused rmsnorm / rope /
silu-mul), used by bench.py, smoke() and the parity tests the way the reference's
bench_one_batch.py drives ModelRunner (python/sglang/bench_one_batch.py:138-268).

Mirrors LlamaDecoderLayer.forward / LlamaModel.forward (reference python/sglang/srt/models/llama.py:
94-98 MLP, :118-190 attention, :245-330 layer + model loop) and Qwen2's qkv bias (qwen2.py:138-143).
It is a measurement harness, not a model zoo: weights are seeded synthetic tensors pushed through
the layers' own TP-aware weight loaders.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Optional

import torch
from torch import nn

from ... import kernels
from ..distributed.parallel_state import (get_tensor_model_parallel_rank, get_tensor_model_parallel_world_size,
                                          get_tp_comm, tensor_model_parallel_all_gather,
                                          tensor_model_parallel_all_reduce_max)
from ..layers.linear import MergedColumnParallelLinear, QKVParallelLinear, RowParallelLinear
from ..layers.radix_attention import RadixAttention


@dataclass
class StackConfig:
    name: str
    hidden_size: int
    num_layers: int
    num_attention_heads: int
    num_kv_heads: int
    head_dim: int
    intermediate_size: int
    vocab_size: int
    rope_theta: float = 500000.0
    rms_norm_eps: float = 1e-5
    qkv_bias: bool = False
    context_len: int = 4096
    is_encoder_decoder: bool = False
    tie_word_embeddings: bool = False  # lm_head shares the embedding matrix (HF config of Qwen2.5-0.5B; qwen2.py:419-421)

    def get_num_kv_heads(self, tp: int) -> int:
        return max(1, self.num_kv_heads // tp)


# shapes from the reference's own benchmark table (sgl-kernel/benchmark/bench_int8_gemm.py:15-58) + HF configs
CONFIGS = {
    "qwen2.5-0.5b": StackConfig("qwen2.5-0.5b", 896, 24, 14, 2, 64, 4864, 151936, 1e6, 1e-6, True,
                                tie_word_embeddings=True),
    "llama-3.1-8b": StackConfig("llama-3.1-8b", 4096, 32, 32, 8, 128, 14336, 128256, 5e5, 1e-5, False),
    "qwen2.5-32b": StackConfig("qwen2.5-32b", 5120, 64, 40, 8, 128, 27648, 152064, 1e6, 1e-6, True),
    "llama-3.1-70b": StackConfig("llama-3.1-70b", 8192, 80, 64, 8, 128, 28672, 128256, 5e5, 1e-5, False),
}


def compute_cos_sin_cache(head_dim: int, max_pos: int, base: float) -> torch.Tensor:
    """RotaryEmbedding._compute_cos_sin_cache (rotary_embedding.py:115-137), fp32."""
    inv_freq = 1.0 / (base ** (torch.arange(0, head_dim, 2, dtype=torch.float) / head_dim))
    t = torch.arange(max_pos, dtype=torch.float)
    freqs = torch.einsum("i,j -> ij", t, inv_freq)
    return torch.cat((freqs.cos(), freqs.sin()), dim=-1)


class DecoderLayer(nn.Module):
    def __init__(self, cfg: StackConfig, layer_id: int, quant_config, dtype):
        super().__init__()
        tp = get_tensor_model_parallel_world_size()
        self.cfg = cfg
        self.qkv_proj = QKVParallelLinear(cfg.hidden_size, cfg.head_dim, cfg.num_attention_heads, cfg.num_kv_heads,
                                          bias=cfg.qkv_bias, quant_config=quant_config, params_dtype=dtype,
                                          prefix=f"layers.{layer_id}.qkv_proj")
        self.o_proj = RowParallelLinear(cfg.num_attention_heads * cfg.head_dim, cfg.hidden_size, bias=False,
                                        quant_config=quant_config, params_dtype=dtype,
                                        prefix=f"layers.{layer_id}.o_proj")
        self.gate_up_proj = MergedColumnParallelLinear(cfg.hidden_size, [cfg.intermediate_size] * 2, bias=False,
                                                       quant_config=quant_config, params_dtype=dtype,
                                                       prefix=f"layers.{layer_id}.gate_up_proj")
        self.down_proj = RowParallelLinear(cfg.intermediate_size, cfg.hidden_size, bias=False,
                                           quant_config=quant_config, params_dtype=dtype,
                                           prefix=f"layers.{layer_id}.down_proj")
        self.num_heads = cfg.num_attention_heads // tp
        self.num_kv_heads = cfg.get_num_kv_heads(tp)
        self.q_size = self.num_heads * cfg.head_dim
        self.kv_size = self.num_kv_heads * cfg.head_dim
        self.attn = RadixAttention(self.num_heads, cfg.head_dim, cfg.head_dim ** -0.5, self.num_kv_heads, layer_id)
        import os

        self.fused_int8 = (type(self.qkv_proj.quant_method).__name__ == "B200W8A8Int8LinearMethod"
                           and os.environ.get("B200_FUSE_STEPS", "1") != "0")
        self.fuse_attn_quant = os.environ.get("B200_FUSE_ATTN_QUANT", "1") != "0"  # A/B switch
        # block-scaled FP8 ([128, 128]) linears: norm + group quant and silu-mul + group quant in one launch each
        self.fused_fp8 = (type(self.qkv_proj.quant_method).__name__ == "B200Fp8LinearMethod"
                          and os.environ.get("B200_FUSE_STEPS", "1") != "0" and cfg.hidden_size % 128 == 0
                          and (cfg.intermediate_size // tp) % 128 == 0)
        self.input_layernorm = nn.Parameter(torch.ones(cfg.hidden_size, dtype=dtype), requires_grad=False)
        self.post_attention_layernorm = nn.Parameter(torch.ones(cfg.hidden_size, dtype=dtype), requires_grad=False)

    def forward(self, positions, hidden_states, forward_batch, residual, cos_sin_cache, comm=None):
        if self.fused_int8 and comm is not None:
            return self.forward_fused_int8_tp(positions, hidden_states, forward_batch, residual, cos_sin_cache, comm)
        if self.fused_int8:
            return self.forward_fused_int8(positions, hidden_states, forward_batch, residual, cos_sin_cache)
        if comm is not None:
            return self.forward_fused_tp(positions, hidden_states, forward_batch, residual, cos_sin_cache, comm)
        if self.fused_fp8:
            return self.forward_fused_fp8(positions, hidden_states, forward_batch, residual, cos_sin_cache)
        eps = self.cfg.rms_norm_eps
        if residual is None:
            residual = hidden_states
            hidden_states = kernels.rmsnorm(hidden_states, self.input_layernorm, eps)
        else:
            kernels.fused_add_rmsnorm(hidden_states, residual, self.input_layernorm, eps)
        qkv, _ = self.qkv_proj(hidden_states)
        q, k, v = qkv.split([self.q_size, self.kv_size, self.kv_size], dim=-1)
        pool = forward_batch.token_to_kv_pool
        cast_kv = pool.dtype != q.dtype  # FP8 KV cache: rope only here, the backend's cast-and-scatter writes the cache
        kernels.rope_neox_set_kv(positions, q, k, v, self.cfg.head_dim, cos_sin_cache,
                                 None if cast_kv else pool.get_key_buffer(self.attn.layer_id),
                                 None if cast_kv else pool.get_value_buffer(self.attn.layer_id),
                                 None if cast_kv else forward_batch.out_cache_loc)
        attn_out = self.attn(q, k, v, forward_batch, save_kv_cache=cast_kv, kv_in_cache=not cast_kv)
        hidden_states, _ = self.o_proj(attn_out)
        kernels.fused_add_rmsnorm(hidden_states, residual, self.post_attention_layernorm, eps)
        gate_up, _ = self.gate_up_proj(hidden_states)
        act = kernels.silu_and_mul(gate_up)
        hidden_states, _ = self.down_proj(act)
        return hidden_states, residual


    def forward_fused_fp8(self, positions, hidden_states, forward_batch, residual, cos_sin_cache):
        """Block-FP8 layer with the activation quantisers fused behind their producers (bit-identical to the un-fused
        pairs): 11 launches per layer instead of 15."""
        eps = self.cfg.rms_norm_eps
        if residual is None:
            residual = hidden_states
            act = kernels.rmsnorm_group_quant_fp8(hidden_states, self.input_layernorm, eps)
        else:
            act = kernels.rmsnorm_group_quant_fp8(hidden_states, self.input_layernorm, eps, residual=residual)
        qkv, _ = self.qkv_proj(act)
        q, k, v = qkv.split([self.q_size, self.kv_size, self.kv_size], dim=-1)
        pool = forward_batch.token_to_kv_pool
        cast_kv = pool.dtype != q.dtype
        kernels.rope_neox_set_kv(positions, q, k, v, self.cfg.head_dim, cos_sin_cache,
                                 None if cast_kv else pool.get_key_buffer(self.attn.layer_id),
                                 None if cast_kv else pool.get_value_buffer(self.attn.layer_id),
                                 None if cast_kv else forward_batch.out_cache_loc)
        attn_out = self.attn(q, k, v, forward_batch, save_kv_cache=cast_kv, kv_in_cache=not cast_kv)
        hidden_states, _ = self.o_proj(attn_out)
        act = kernels.rmsnorm_group_quant_fp8(hidden_states, self.post_attention_layernorm, eps, residual=residual)
        gate_up, _ = self.gate_up_proj(act)
        hidden_states, _ = self.down_proj(kernels.silu_and_mul_group_quant_fp8(gate_up))
        return hidden_states, residual

    def _forward_fused_int8(self, positions, hidden_states, forward_batch, residual, cos_sin_cache):
        """Same layer with the §8f #2 fusions: norm+quant, rope+KV-write, silu-mul+quant (10 launches, not 15)."""
        eps = self.cfg.rms_norm_eps
        if residual is None:
            residual = hidden_states
            act = kernels.rmsnorm_quant_int8(hidden_states, self.input_layernorm, eps)
        else:
            act = kernels.rmsnorm_quant_int8(hidden_states, self.input_layernorm, eps, residual=residual)
        qkv, _ = self.qkv_proj(act)
        q, k, v = qkv.split([self.q_size, self.kv_size, self.kv_size], dim=-1)
        pool = forward_batch.token_to_kv_pool
        cast_kv = pool.dtype != q.dtype  # FP8 KV cache: rope only here, the backend's cast-and-scatter writes the cache
        kernels.rope_neox_set_kv(positions, q, k, v, self.cfg.head_dim, cos_sin_cache,
                                 None if cast_kv else pool.get_key_buffer(self.attn.layer_id),
                                 None if cast_kv else pool.get_value_buffer(self.attn.layer_id),
                                 None if cast_kv else forward_batch.out_cache_loc)
        attn_out = self.attn(q, k, v, forward_batch, save_kv_cache=cast_kv, int8_out=self.fuse_attn_quant,
                             kv_in_cache=not cast_kv)
        hidden_states, _ = self.o_proj(attn_out)
        act = kernels.rmsnorm_quant_int8(hidden_states, self.post_attention_layernorm, eps, residual=residual)
        gate_up, _ = self.gate_up_proj(act)
        hidden_states, _ = self.down_proj(kernels.silu_and_mul_quant_int8(gate_up))
        return hidden_states, residual

    forward_fused_int8 = _forward_fused_int8

    def forward_fused_int8_tp(self, positions, hidden_states, forward_batch, residual, cos_sin_cache, comm):
        """TP > 1 with the fused exchange: o_proj / down_proj write their partial sums into the
        symmetric buffer and the NEXT norm is the exchange kernel (reduce over ranks + residual + RMSNorm + int8 quant
        + all-gather of the int8 rows).  `hidden_states` arrives un-reduced except at the first layer (residual None);
        it leaves un-reduced, to be consumed by the next layer's input norm or the stack's final norm."""
        eps = self.cfg.rms_norm_eps
        M = hidden_states.shape[0]
        if residual is None:
            residual = hidden_states
            act = kernels.rmsnorm_quant_int8(hidden_states, self.input_layernorm, eps)
        else:
            act = comm.reduce_rmsnorm_quant_int8(M, self.input_layernorm, eps, residual)
        qkv, _ = self.qkv_proj(act)
        q, k, v = qkv.split([self.q_size, self.kv_size, self.kv_size], dim=-1)
        pool = forward_batch.token_to_kv_pool
        cast_kv = pool.dtype != q.dtype  # FP8 KV cache: rope only here, the backend's cast-and-scatter writes the cache
        kernels.rope_neox_set_kv(positions, q, k, v, self.cfg.head_dim, cos_sin_cache,
                                 None if cast_kv else pool.get_key_buffer(self.attn.layer_id),
                                 None if cast_kv else pool.get_value_buffer(self.attn.layer_id),
                                 None if cast_kv else forward_batch.out_cache_loc)
        attn_out = self.attn(q, k, v, forward_batch, save_kv_cache=cast_kv, int8_out=self.fuse_attn_quant,
                             kv_in_cache=not cast_kv)
        self.o_proj(attn_out, can_fuse_mlp_allreduce=True, out=comm.partial(M))
        act = comm.reduce_rmsnorm_quant_int8(M, self.post_attention_layernorm, eps, residual)
        gate_up, _ = self.gate_up_proj(act)
        hidden_states, _ = self.down_proj(kernels.silu_and_mul_quant_int8(gate_up), can_fuse_mlp_allreduce=True,
                                          out=comm.partial(M))
        return hidden_states, residual


    def forward_fused_tp(self, positions, hidden_states, forward_batch, residual, cos_sin_cache, comm):
        """TP > 1, FP8 / AWQ / un-quantised linears, with the fused exchange: the same protocol as
        forward_fused_int8_tp, but the exchange kernel hands every rank the normalised rows in the activation dtype
        (the reference's allreduce + rmsnorm fusion, layers/flashinfer_comm_fusion.py) and the linear method quantises
        them itself.  Replaces dist.all_reduce + fused_add_rmsnorm per row-parallel linear."""
        eps = self.cfg.rms_norm_eps
        M = hidden_states.shape[0]
        if residual is None:
            residual = hidden_states
            act = kernels.rmsnorm(hidden_states, self.input_layernorm, eps)
        else:
            act = comm.reduce_rmsnorm(M, self.input_layernorm, eps, residual)
        qkv, _ = self.qkv_proj(act)
        q, k, v = qkv.split([self.q_size, self.kv_size, self.kv_size], dim=-1)
        pool = forward_batch.token_to_kv_pool
        cast_kv = pool.dtype != q.dtype
        kernels.rope_neox_set_kv(positions, q, k, v, self.cfg.head_dim, cos_sin_cache,
                                 None if cast_kv else pool.get_key_buffer(self.attn.layer_id),
                                 None if cast_kv else pool.get_value_buffer(self.attn.layer_id),
                                 None if cast_kv else forward_batch.out_cache_loc)
        attn_out = self.attn(q, k, v, forward_batch, save_kv_cache=cast_kv, kv_in_cache=not cast_kv)
        self.o_proj(attn_out, can_fuse_mlp_allreduce=True, out=comm.partial(M))
        act = comm.reduce_rmsnorm(M, self.post_attention_layernorm, eps, residual)
        gate_up, _ = self.gate_up_proj(act)
        hidden_states, _ = self.down_proj(kernels.silu_and_mul(gate_up), can_fuse_mlp_allreduce=True, out=comm.partial(M))
        return hidden_states, residual


class LayerStack(nn.Module):
    """embed -> N decoder layers -> final norm -> lm_head (vocab-parallel) -> greedy argmax."""

    def __init__(self, cfg: StackConfig, quant_config, dtype=torch.bfloat16, num_layers: Optional[int] = None,
                 max_positions: Optional[int] = None):
        super().__init__()
        self.cfg = cfg
        self.dtype = dtype
        tp = get_tensor_model_parallel_world_size()
        self.tp_size = tp
        n_layers = num_layers if num_layers is not None else cfg.num_layers
        self.layers = nn.ModuleList([DecoderLayer(cfg, i, quant_config, dtype) for i in range(n_layers)])
        # the embedding is replicated (a gather, not on the hot path); lm_head is vocab-parallel and its
        # logits are all-gathered (logits_processor.py), as in the reference
        self.vocab_per_rank = (cfg.vocab_size + tp - 1) // tp
        self.embed_tokens = nn.Parameter(torch.empty(cfg.vocab_size, cfg.hidden_size, dtype=dtype), requires_grad=False)
        self.lm_head = nn.Parameter(torch.empty(self.vocab_per_rank, cfg.hidden_size, dtype=dtype), requires_grad=False)
        self.norm = nn.Parameter(torch.ones(cfg.hidden_size, dtype=dtype), requires_grad=False)
        self.register_buffer("cos_sin_cache",
                             compute_cos_sin_cache(cfg.head_dim, max_positions or cfg.context_len + 8, cfg.rope_theta),
                             persistent=False)

    def forward(self, input_ids, positions, forward_batch, seq_lens_for_positions=None):
        """`seq_lens_for_positions` (decode): the embedding launch also writes positions = seq_lens - 1
        (forward_batch_info.py:960-962) into `positions`."""
        hidden_states = kernels.embedding(input_ids, self.embed_tokens, seq_lens_for_positions,
                                          positions if seq_lens_for_positions is not None else None)
        residual = None
        comm = get_tp_comm() if (self.tp_size > 1 and len(self.layers) > 0) else None
        if comm is not None and hidden_states.shape[0] > comm.max_tokens:
            comm = None  # larger than the symmetric buffers: NCCL all-reduce path
        for layer in self.layers:
            hidden_states, residual = layer(positions, hidden_states, forward_batch, residual, self.cos_sin_cache, comm)
        if comm is not None and len(self.layers) > 0:
            # final norm = the last exchange: every rank receives the normalised rows of all tokens
            return comm.reduce_rmsnorm(hidden_states.shape[0], self.norm, self.cfg.rms_norm_eps, residual)
        kernels.fused_add_rmsnorm(hidden_states, residual, self.norm, self.cfg.rms_norm_eps)
        return hidden_states

    def local_logits(self, hidden_states):
        """This rank's [M, vocab_per_rank] slice of the logits (logits_processor.py:470-500).  Decode-size M runs the
        repo's own weight-streaming tcgen05 GEMM; a larger M (every position of a prefill chunk — not something the
        greedy path asks for) is a plain library GEMM."""
        if hidden_states.shape[0] <= 256:
            return kernels.dense_gemm_nt(hidden_states.contiguous(), self.lm_head)
        return torch.matmul(hidden_states, self.lm_head.t())

    def logits(self, hidden_states):
        """Full logits: vocab-parallel lm_head + all-gather (what a non-greedy sampler needs)."""
        local = self.local_logits(hidden_states)
        if self.tp_size > 1:
            local = tensor_model_parallel_all_gather(local, dim=-1)
        return local[..., : self.cfg.vocab_size]

    def greedy(self, hidden_states, out=None, key_buf=None):
        """Greedy next-token ids (sampler.py: argmax over the logits).  At TP > 1 the logits are never gathered: each
        rank reduces its vocabulary slice to one 8-byte (value, global index) key per row and the ranks max-combine
        the keys — 8 B/row on the wire instead of 2 * vocab B/row."""
        local = self.local_logits(hidden_states)
        if self.tp_size == 1:
            return kernels.argmax(local[:, : self.cfg.vocab_size], out=out)
        rank = get_tensor_model_parallel_rank()
        valid = max(0, min(self.vocab_per_rank, self.cfg.vocab_size - rank * self.vocab_per_rank))
        assert valid > 0, "vocab-parallel argmax: a rank without vocabulary rows"
        keys = kernels.argmax_key(local[:, :valid], rank * self.vocab_per_rank, out=key_buf)
        tensor_model_parallel_all_reduce_max(keys)
        return kernels.argmax_key_decode(keys, out=out)


# ---- seeded synthetic weights through the layers' own loaders ----------------------------------------
def _quant_name(layer) -> str:
    return type(layer.quant_method).__name__


def _gen(shape, gen, device, kind, k_dim=None, gain=1.0):
    if kind == "int8":
        return torch.randint(-127, 128, shape, generator=gen, device=device, dtype=torch.int32).to(torch.int8)
    if kind == "scale_int8":  # effective weight std ~ gain / sqrt(K)
        return (torch.rand(shape, generator=gen, device=device) + 0.5) * (gain / (73.3 * math.sqrt(k_dim)))
    if kind == "dense":
        return torch.randn(shape, generator=gen, device=device) * (gain / math.sqrt(k_dim))
    raise ValueError(kind)


def load_synthetic_weights(stack: LayerStack, seed: int = 1234, device="cuda", gain: float = 1.0):
    """Fill every parameter from a seeded generator.  FULL (un-sharded) logical tensors are generated and
    handed to the layers' weight loaders, which shard them per TP rank — identical tensors on every rank
    (same seed), as when loading a checkpoint."""
    gen = torch.Generator(device=device).manual_seed(seed)
    cfg = stack.cfg
    dt = stack.dtype

    def fill_linear(lin, out_sizes, shard_ids, in_size):
        method = _quant_name(lin)
        for sid, n in zip(shard_ids, out_sizes):
            if method == "B200W8A8Int8LinearMethod":
                w = _gen((n, in_size), gen, device, "int8")
                s = _gen((n, 1), gen, device, "scale_int8", in_size, gain)
                for p, t in ((lin.weight, w), (lin.weight_scale, s)):
                    lin.weight_loader(p, t, sid) if sid is not None else lin.weight_loader(p, t)
            elif method == "UnquantizedLinearMethod":
                w = _gen((n, in_size), gen, device, "dense", in_size, gain).to(dt)
                lin.weight_loader(lin.weight, w, sid) if sid is not None else lin.weight_loader(lin.weight, w)
            else:
                from ..layers.quantization import synthetic  # fp8 / awq generators live with their methods

                synthetic.fill(lin, sid, n, in_size, gen, device, gain)
        if getattr(lin, "bias", None) is not None:
            lin.bias.data.copy_((torch.randn(lin.bias.shape, generator=gen, device=device) * 0.1).to(dt))

    for layer in stack.layers:
        hq, hkv, d = cfg.num_attention_heads, cfg.num_kv_heads, cfg.head_dim
        fill_linear(layer.qkv_proj, [hq * d, hkv * d, hkv * d], ["q", "k", "v"], cfg.hidden_size)
        fill_linear(layer.o_proj, [cfg.hidden_size], [None], hq * d)
        fill_linear(layer.gate_up_proj, [cfg.intermediate_size] * 2, [0, 1], cfg.hidden_size)
        fill_linear(layer.down_proj, [cfg.hidden_size], [None], cfg.intermediate_size)
        for norm in (layer.input_layernorm, layer.post_attention_layernorm):
            norm.data.copy_((1.0 + 0.1 * torch.randn(norm.shape, generator=gen, device=device)).to(dt))
        layer.qkv_proj.quant_method.process_weights_after_loading(layer.qkv_proj)
        layer.o_proj.quant_method.process_weights_after_loading(layer.o_proj)
        layer.gate_up_proj.quant_method.process_weights_after_loading(layer.gate_up_proj)
        layer.down_proj.quant_method.process_weights_after_loading(layer.down_proj)
    stack.norm.data.copy_((1.0 + 0.1 * torch.randn(stack.norm.shape, generator=gen, device=device)).to(dt))
    # embeddings / lm_head in row chunks (the full fp32 randn of a 128k x 4k matrix would be 2 GiB)
    rank = get_tensor_model_parallel_rank()
    for name, rows0, rows in (("embed_tokens", 0, cfg.vocab_size),
                              ("lm_head", rank * stack.vocab_per_rank, stack.vocab_per_rank)):
        p = getattr(stack, name)
        g2 = torch.Generator(device=device).manual_seed(seed + (1 if name == "lm_head" else 2))
        chunk = 8192
        # generate ALL rows in order so every rank sees the same logical matrix, keep only its slice
        total = cfg.vocab_size if name == "embed_tokens" else stack.vocab_per_rank * stack.tp_size
        for r in range(0, total, chunk):
            n = min(chunk, total - r)
            blk = (torch.randn((n, cfg.hidden_size), generator=g2, device=device) * (gain / math.sqrt(cfg.hidden_size))
                   if name == "lm_head" else torch.randn((n, cfg.hidden_size), generator=g2, device=device))
            lo, hi = max(r, rows0), min(r + n, rows0 + rows)
            if lo < hi:
                p.data[lo - rows0: hi - rows0].copy_(blk[lo - r: hi - r].to(dt))
    if cfg.tie_word_embeddings:  # this rank's vocabulary slice of the embedding matrix (rows past the vocabulary: zero)
        lo = rank * stack.vocab_per_rank
        hi = min(lo + stack.vocab_per_rank, cfg.vocab_size)
        stack.lm_head.data.zero_()
        if hi > lo:
            stack.lm_head.data[: hi - lo].copy_(stack.embed_tokens.data[lo:hi])
