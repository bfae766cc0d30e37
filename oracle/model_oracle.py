"""TEST INFRASTRUCTURE (oracle) — the decoder-layer loop on CPU, built from the oracle restatements.

Restates LlamaDecoderLayer.forward / LlamaModel.forward (reference python/sglang/srt/models/llama.py:
245-330; MLP :94-98; attention :160-190) with torch_native attention (attention_oracle) and the
quantised-linear restatements (quant_oracle).  Weights are taken from a device LayerStack so both sides
hold identical tensors.  Parity: pinned through its parts (see attention_oracle / quant_oracle headers).
"""
from __future__ import annotations

import torch

from . import attention_oracle as ao
from . import quant_oracle as qo


def _cpu(t):
    return None if t is None else t.detach().to("cpu")


class OracleLinear:
    def __init__(self, lin):
        self.kind = type(lin.quant_method).__name__
        self.bias = _cpu(getattr(lin, "bias", None))
        if self.kind == "B200W8A8Int8LinearMethod":
            self.w_q = _cpu(lin.weight).t().contiguous()  # back to [N, K]
            self.w_s = _cpu(lin.weight_scale)
        elif self.kind == "B200Fp8LinearMethod":
            self.w_q = _cpu(lin.weight)
            self.w_s = _cpu(lin.weight_scale_inv)
        elif self.kind == "B200AWQLinearMethod":
            self.qweight, self.scales, self.qzeros = _cpu(lin.qweight), _cpu(lin.scales), _cpu(lin.qzeros)
            self.group = lin.quant_method.quant_config.group_size
        else:
            self.w = _cpu(lin.weight)

    def __call__(self, x, add_bias=True):
        bias = self.bias if add_bias else None
        if self.kind == "B200W8A8Int8LinearMethod":
            return qo.w8a8_int8_linear(x, self.w_q, self.w_s, bias)
        if self.kind == "B200Fp8LinearMethod":
            return qo.w8a8_block_fp8_linear(x, self.w_q, self.w_s, bias)
        if self.kind == "B200AWQLinearMethod":
            return qo.awq_linear(x, self.qweight, self.scales, self.qzeros, self.group, bias)
        return torch.nn.functional.linear(x.float(), self.w.float(), None if bias is None else bias.float()).to(x.dtype)


class OracleStack:
    """CPU twin of srt/model_executor/layer_stack.LayerStack (TP = 1)."""

    def __init__(self, stack):
        self.cfg = stack.cfg
        self.layers = []
        for l in stack.layers:
            self.layers.append(dict(qkv=OracleLinear(l.qkv_proj), o=OracleLinear(l.o_proj),
                                    gate_up=OracleLinear(l.gate_up_proj), down=OracleLinear(l.down_proj),
                                    ln1=_cpu(l.input_layernorm), ln2=_cpu(l.post_attention_layernorm),
                                    q_size=l.q_size, kv_size=l.kv_size))
        self.embed = _cpu(stack.embed_tokens)
        self.lm_head = _cpu(stack.lm_head)
        self.norm = _cpu(stack.norm)
        self.cos_sin = _cpu(stack.cos_sin_cache)

    def forward(self, input_ids, positions, k_caches, v_caches, req_to_token, req_pool_indices, seq_lens,
                out_cache_loc, prefix_lens=None, record=None):
        """Decode when prefix_lens is None, else extend.  k/v_caches: per-layer CPU pools (updated in place).
        record: optional list that receives, per layer, (hidden_in, residual_in, hidden_out, residual_out) — the
        teacher-forcing points of the per-layer parity test."""
        cfg = self.cfg
        eps = cfg.rms_norm_eps
        h = self.embed[input_ids]
        residual = None
        hq, hkv, d = cfg.num_attention_heads, cfg.num_kv_heads, cfg.head_dim
        for i, L in enumerate(self.layers):
            if record is not None:
                if i > 0:
                    record[-1] += (h.clone(), residual.clone())
                record.append((h.clone(), None if residual is None else residual.clone()))
            if residual is None:
                residual = h
                h = qo.rmsnorm(h, L["ln1"], eps)
            else:
                h, residual = qo.rmsnorm(h, L["ln1"], eps, residual)
            qkv = L["qkv"](h)
            q, k, v = qkv.split([L["q_size"], L["kv_size"], L["kv_size"]], dim=-1)
            q, k = qo.rope_neox(positions, q, k, d, self.cos_sin)
            T = q.shape[0]
            kn, vn = k.reshape(T, hkv, d), v.reshape(T, hkv, d)
            if k_caches[i].dtype == torch.uint8:  # --kv-cache-dtype fp8_e4m3 (stored as uint8, memory_pool.py:114-118)
                if prefix_lens is None:
                    ao.set_kv_buffer_fp8(k_caches[i], v_caches[i], out_cache_loc, kn, vn)
                    o = ao.decode_attention_fp8kv(q.reshape(T, hq, d), k_caches[i], v_caches[i], req_to_token,
                                                  req_pool_indices, seq_lens, d ** -0.5)
                else:
                    ext = (seq_lens - prefix_lens).to(torch.int32)
                    o = ao.extend_attention_fp8kv(q.reshape(T, hq, d), kn, vn, k_caches[i], v_caches[i], req_to_token,
                                                  req_pool_indices, seq_lens, prefix_lens, ext, out_cache_loc, d ** -0.5)
                    ao.set_kv_buffer_fp8(k_caches[i], v_caches[i], out_cache_loc, kn, vn)
                h = L["o"](o.reshape(T, hq * d))
                h, residual = qo.rmsnorm(h, L["ln2"], eps, residual)
                h = L["down"](qo.silu_and_mul(L["gate_up"](h)))
                continue
            ao.set_kv_buffer(k_caches[i], v_caches[i], out_cache_loc, kn, vn)
            if prefix_lens is None:
                o = ao.decode_attention(q.reshape(T, hq, d), k_caches[i], v_caches[i], req_to_token, req_pool_indices,
                                        seq_lens, d ** -0.5)
            else:
                ext = (seq_lens - prefix_lens).to(torch.int32)
                o = ao.extend_attention(q.reshape(T, hq, d), k_caches[i], v_caches[i], req_to_token, req_pool_indices,
                                        seq_lens, prefix_lens, ext, d ** -0.5)
            h = L["o"](o.reshape(T, hq * d))
            h, residual = qo.rmsnorm(h, L["ln2"], eps, residual)
            gu = L["gate_up"](h)
            h = L["down"](qo.silu_and_mul(gu))
        if record is not None and self.layers:
            record[-1] += (h.clone(), residual.clone())
        h, _ = qo.rmsnorm(h, self.norm, eps, residual)
        return h

    def logits(self, h):
        return (h.float() @ self.lm_head.float().t()).to(h.dtype)[..., : self.cfg.vocab_size]
