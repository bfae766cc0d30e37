"""CPU: pin the oracle restatement against the reference's OWN compiled CPU kernels (oracle/_ref,
built from /root/reference/sgl-kernel/csrc/cpu by oracle/build_ref.py), with the reference tests'
shapes and tolerances (test/srt/cpu/test_decode.py:62-143, test_extend.py:78-183, test_gemm.py:83-126)."""
import pytest
import torch

from oracle import attention_oracle as ao
from oracle import build_ref
from oracle import quant_oracle as qo
from sglang_npu_b200 import synth


@pytest.fixture(scope="module")
def ref_ops():
    build_ref.build()
    ops = build_ref.load()
    if ops is None:
        pytest.skip("oracle/_ref not available on this host (needs AVX512-BF16 + AMX)")
    return ops


@pytest.mark.parametrize("M,N,K", [(1, 32 * 12, 32 * 17), (8, 512, 544), (64, 1152, 896), (101, 896, 4864)])
@pytest.mark.parametrize("has_bias", [False, True])
def test_int8_linear_vs_reference_cpu_kernel(ref_ops, M, N, K, has_bias):
    torch.manual_seed(1234)
    x = synth.activations(M, K, torch.bfloat16, seed=M)
    w, ws = synth.int8_weight(N, K, seed=N)
    bias = torch.randn(N) if has_bias else None
    packed = ref_ops.convert_weight_packed(w)
    # (1) the quantisation step: the AVX512 kernel (cpu/vec.h:133-188) stores u8 = q + 128 and may round a
    #     handful of near-tie values the other way; the oracle fixes torch.round's rule.
    xq_ref, xs_ref = ref_ops.per_token_quant_int8_cpu(x)
    xq, xs = qo.per_token_quant_int8(x)
    codes_ref = (xq_ref.to(torch.int16) - 128)
    diff = (codes_ref - xq.to(torch.int16)).abs()
    assert diff.max() <= 1 and (diff != 0).float().mean() < 1e-3
    torch.testing.assert_close(xs_ref.view(-1), xs.view(-1), rtol=1e-6, atol=0)
    # (2) GEMM + epilogue on IDENTICAL codes: the restatement must agree to one bf16 rounding
    out_mm = ref_ops.int8_scaled_mm_cpu(xq_ref, packed, xs_ref.view(-1), ws.view(-1), bias, torch.bfloat16, True)
    ref_mm = qo.int8_scaled_mm(codes_ref.to(torch.int8), w, xs_ref.view(-1), ws, torch.bfloat16, bias)
    ulp = torch.maximum(ref_mm.float().abs(), torch.tensor(1e-3)) * 2.0 ** -7
    assert ((out_mm.float() - ref_mm.float()).abs() <= ulp).all()
    # (3) fused quant + GEMM, the op W8A8Int8LinearMethod.apply calls on CPU (w8a8_int8.py:381-388): same
    #     result up to the few differing activation codes — cosine / max-abs criterion of north_star
    out = ref_ops.int8_scaled_mm_with_quant(x, packed, ws.view(-1), bias, torch.bfloat16, True)
    ref = qo.w8a8_int8_linear(x, w, ws, bias)
    cos = torch.nn.functional.cosine_similarity(out.flatten().float(), ref.flatten().float(), dim=0)
    assert cos.item() > 0.9999
    assert ((out.float() - ref.float()).abs() <= 2 * ulp + 2e-2).all()


@pytest.mark.parametrize("B,HQ,HKV,D", [(1, 32, 8, 128), (3, 40, 8, 128), (2, 14, 2, 64)])
def test_decode_attention_vs_reference_cpu_kernel(ref_ops, B, HQ, HKV, D):
    lens = [257, 1024, 33][:B]
    pb = synth.make_paged_batch(lens, 16, seed=B)
    k, v = synth.randn_kv(pb.num_slots, 16, HKV, D, torch.bfloat16, seed=3)
    g = torch.Generator().manual_seed(5)
    q = torch.randn((B, HQ, D), generator=g).bfloat16()
    key = torch.randn((B, HKV, D), generator=g).bfloat16()
    val = torch.randn((B, HKV, D), generator=g).bfloat16()
    loc = torch.stack([pb.req_to_token[pb.req_pool_indices[i], lens[i] - 1] for i in range(B)]).long()
    k1, v1 = k.clone(), v.clone()
    o = torch.zeros_like(q)
    attn_logits = torch.zeros((B, HQ, 8, D + 1), dtype=torch.float32)  # zeros, as intel_amx_backend.py:36-45 allocates it
    ref_ops.decode_attention_cpu(q, k1, v1, o, key, val, loc, attn_logits, pb.req_to_token, pb.req_pool_indices,
                                 pb.seq_lens, D ** -0.5, 0.0)
    ao.set_kv_buffer(k, v, loc, key, val)
    assert torch.equal(k.view(torch.int16), k1.view(torch.int16))  # decode_set_kv_buffer (decode.cpp:771) is a copy
    assert torch.equal(v.view(torch.int16), v1.view(torch.int16))
    ref = ao.decode_attention(q, k, v, pb.req_to_token, pb.req_pool_indices, pb.seq_lens, D ** -0.5)
    cos = torch.nn.functional.cosine_similarity(o.flatten().float(), ref.flatten().float(), dim=0)
    assert cos.item() > 0.99
    torch.testing.assert_close(o, ref, atol=3e-2, rtol=1e-6)


def test_extend_attention_vs_reference_cpu_kernel(ref_ops):
    B, HQ, HKV, D = 3, 8, 2, 128
    lens, prefix = [70, 33, 150], [32, 0, 100]
    pb = synth.make_paged_batch(lens, 16, seed=9)
    k, v = synth.randn_kv(pb.num_slots, 16, HKV, D, torch.bfloat16, seed=4)
    prefix_t = torch.tensor(prefix, dtype=torch.int32)
    ext = pb.seq_lens.to(torch.int32) - prefix_t
    T = int(ext.sum())
    g = torch.Generator().manual_seed(6)
    q = torch.randn((T, HQ, D), generator=g).bfloat16()
    ke = torch.randn((T, HKV, D), generator=g).bfloat16()
    ve = torch.randn((T, HKV, D), generator=g).bfloat16()
    loc = torch.cat([pb.req_to_token[pb.req_pool_indices[i], prefix[i]: lens[i]] for i in range(B)]).long()
    ao.set_kv_buffer(k, v, loc, ke, ve)
    start = torch.zeros(B, dtype=torch.int32)  # same dtype as req_to_token (extend.cpp entry checks)
    start[1:] = torch.cumsum(ext[:-1], 0)
    o = torch.empty_like(q)
    ref_ops.extend_attention_cpu(q, ke, ve, o, k, v, pb.req_to_token, pb.req_pool_indices, pb.seq_lens,
                                 ext, start, int(ext.max()), D ** -0.5, 0.0)
    ref = ao.extend_attention(q, k, v, pb.req_to_token, pb.req_pool_indices, pb.seq_lens, prefix_t, ext, D ** -0.5)
    torch.testing.assert_close(o, ref, atol=1e-2, rtol=1e-2)
