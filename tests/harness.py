"""MockModelRunner-style harness (pattern of the reference's backend unit test,
python/sglang/test/attention/test_flashattn_backend.py:15-69, :163-200): real pools, a hand-built
ForwardBatch, and the backend under test."""
import types

import torch

from sglang_npu_b200.srt.layers.attention.b200_backend import B200AttnBackend
from sglang_npu_b200.srt.layers.radix_attention import RadixAttention
from sglang_npu_b200.srt.mem_cache.memory_pool import MHATokenToKVPool, ReqToTokenPool
from sglang_npu_b200.srt.model_executor.forward_batch_info import ForwardBatch, ForwardMode


class MockModelConfig:
    def __init__(self, hq, hkv, ctx):
        self.num_attention_heads = hq
        self._hkv = hkv
        self.context_len = ctx
        self.is_encoder_decoder = False

    def get_num_kv_heads(self, tp):
        return max(1, self._hkv // tp)


def build_runner(device, req_to_token, num_slots, page_size, hq, hkv, d, dtype, layers=1, ctx=None):
    r2t = ReqToTokenPool(req_to_token.shape[0], req_to_token.shape[1], device)
    r2t.req_to_token.copy_(req_to_token)
    kv = MHATokenToKVPool(num_slots, page_size, dtype, hkv, d, layers, device)
    runner = types.SimpleNamespace(
        device=device, req_to_token_pool=r2t, token_to_kv_pool=kv,
        model_config=MockModelConfig(hq, hkv, ctx or req_to_token.shape[1] - 4),
        sliding_window_size=None, tp_size=1, server_args=types.SimpleNamespace(), gpu_id=0)
    backend = B200AttnBackend(runner)
    return runner, backend


def decode_batch(runner, backend, req_pool_indices, seq_lens, out_cache_loc, with_cpu_lens=True,
                 shared_prefix_len=0):
    dev = runner.device
    fb = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=len(seq_lens), input_ids=None,
                      req_pool_indices=req_pool_indices.to(dev), seq_lens=seq_lens.to(dev),
                      out_cache_loc=out_cache_loc.to(dev), seq_lens_sum=int(seq_lens.sum()),
                      seq_lens_cpu=seq_lens.cpu() if with_cpu_lens else None,
                      req_to_token_pool=runner.req_to_token_pool, token_to_kv_pool=runner.token_to_kv_pool,
                      attn_backend=backend, shared_prefix_len=shared_prefix_len)
    backend.init_forward_metadata(fb)
    return fb


def extend_batch(runner, backend, req_pool_indices, seq_lens, prefix_lens, out_cache_loc):
    dev = runner.device
    ext = (seq_lens - prefix_lens.to(seq_lens.dtype)).to(torch.int32)
    fb = ForwardBatch(forward_mode=ForwardMode.EXTEND, batch_size=len(seq_lens), input_ids=None,
                      req_pool_indices=req_pool_indices.to(dev), seq_lens=seq_lens.to(dev),
                      out_cache_loc=out_cache_loc.to(dev), seq_lens_sum=int(seq_lens.sum()),
                      seq_lens_cpu=seq_lens.cpu(), extend_num_tokens=int(ext.sum()),
                      extend_seq_lens=ext.to(dev), extend_prefix_lens=prefix_lens.to(torch.int32).to(dev),
                      extend_prefix_lens_cpu=prefix_lens.tolist(), extend_seq_lens_cpu=ext.tolist(),
                      req_to_token_pool=runner.req_to_token_pool, token_to_kv_pool=runner.token_to_kv_pool,
                      attn_backend=backend)
    backend.init_forward_metadata(fb)
    return fb


def make_layer(hq, hkv, d, layer_id=0, logit_cap=0.0):
    return RadixAttention(num_heads=hq, head_dim=d, scaling=d ** -0.5, num_kv_heads=hkv, layer_id=layer_id,
                          logit_cap=logit_cap)
