"""CPU known-answer tests of oracle/quant_oracle.py (no GPU): hand-computable cases of the rules the reference's
own tests define (test/srt/cpu/utils.py:24-32,:91-115; python/sglang/test/test_block_fp8.py:24-49,:274-330;
sgl-kernel/tests/test_awq_dequant.py:9-61)."""
import torch

from oracle import quant_oracle as qo


def test_per_token_quant_int8_known_answer():
    x = torch.tensor([[1.0, -2.0, 0.5, 127.0], [0.0, 0.0, 0.0, 0.0]])
    q, s = qo.per_token_quant_int8(x)
    assert q.dtype == torch.int8 and s.shape == (2, 1)
    assert q[0].tolist() == [1, -2, 0, 127]  # 0.5 -> round-half-even -> 0
    assert abs(s[0, 0].item() - 1.0) < 1e-7
    assert q[1].tolist() == [0, 0, 0, 0] and s[1, 0].item() > 0  # absmax clamp keeps the scale finite


def test_int8_accumulator_is_exact_integer_matmul():
    g = torch.Generator().manual_seed(0)
    a = torch.randint(-127, 128, (5, 64), generator=g, dtype=torch.int32).to(torch.int8)
    w = torch.randint(-127, 128, (7, 64), generator=g, dtype=torch.int32).to(torch.int8)
    acc = qo.int8_accumulate(a, w)
    assert acc.dtype == torch.int32
    assert torch.equal(acc.long(), a.long() @ w.long().t())


def test_block_fp8_group_quant_known_answer():
    x = torch.zeros(1, 256)
    x[0, 0], x[0, 128] = 448.0, -2.0
    q, s = qo.per_token_group_quant_fp8(x, 128)
    assert s.shape == (1, 2) and abs(s[0, 0].item() - 1.0) < 1e-6 and abs(s[0, 1].item() - 2.0 / 448) < 1e-9
    assert q.dtype == torch.float8_e4m3fn and q.float()[0, 0].item() == 448.0 and q.float()[0, 128].item() == -448.0


def test_block_fp8_matmul_applies_scales_per_128_block():
    M, N, K = 2, 128, 256
    a = torch.ones(M, K).to(torch.float8_e4m3fn)
    w = torch.ones(N, K).to(torch.float8_e4m3fn)
    a_s = torch.tensor([[1.0, 2.0], [3.0, 4.0]])
    w_s = torch.tensor([[0.5, 0.25]])
    out = qo.w8a8_block_fp8_matmul(a, w, a_s, w_s, [128, 128], torch.float32)
    assert torch.allclose(out[0], torch.full((N,), 128 * 1.0 * 0.5 + 128 * 2.0 * 0.25))
    assert torch.allclose(out[1], torch.full((N,), 128 * 3.0 * 0.5 + 128 * 4.0 * 0.25))


def test_awq_nibble_order_known_answer():
    # packed word with nibble j holding value j: logical column i reads nibble ORDER[i] = [0,4,1,5,2,6,3,7]
    word = sum(j << (4 * j) for j in range(8))
    qweight = torch.tensor([[word]], dtype=torch.int64).to(torch.int32)
    qzeros = torch.zeros(1, 1, dtype=torch.int32)
    scales = torch.ones(1, 8, dtype=torch.float16)
    w = qo.awq_dequantize(qweight, scales, qzeros, 1)
    assert w.flatten().tolist() == [0, 4, 1, 5, 2, 6, 3, 7]


def test_rmsnorm_silu_rope_shapes():
    x = torch.randn(3, 16)
    assert qo.rmsnorm(x, torch.ones(16), 1e-6).shape == x.shape
    assert qo.silu_and_mul(x).shape == (3, 8)


# ---- FP8 KV cache / shared-prefix host logic (no GPU) ------------------------------------------------------------------
def test_fp8_kv_oracles_are_consistent_with_the_16bit_oracles():
    """decode / extend over an e4m3 cache == the 16-bit oracles run on the up-cast cache (the up-cast is exact)."""
    import torch
    from oracle import attention_oracle as ao
    from sglang_npu_b200 import synth

    hq, hkv, d, page = 8, 2, 64, 16
    lens = [40, 17, 33]
    pb = synth.make_paged_batch(lens, page, seed=4)
    k0, v0 = synth.randn_kv(pb.num_slots, page, hkv, d, torch.bfloat16, seed=5)
    k8 = k0.to(torch.float8_e4m3fn).view(torch.uint8)
    v8 = v0.to(torch.float8_e4m3fn).view(torch.uint8)
    g = torch.Generator().manual_seed(6)
    q = torch.randn((3, hq, d), generator=g).to(torch.bfloat16)
    o8 = ao.decode_attention_fp8kv(q, k8, v8, pb.req_to_token, pb.req_pool_indices, pb.seq_lens, d ** -0.5, 0.5, 2.0)
    kf = k8.view(torch.float8_e4m3fn).float() * 0.5
    vf = v8.view(torch.float8_e4m3fn).float() * 2.0
    o16 = ao.decode_attention(q, kf, vf, pb.req_to_token, pb.req_pool_indices, pb.seq_lens, d ** -0.5,
                              compute_dtype=torch.float32)
    assert torch.equal(o8.view(torch.int16), o16.to(torch.bfloat16).view(torch.int16))
    # the KV write of values that are already e4m3-representable is the identity
    loc = torch.arange(page, page + 5)
    kc, vc = torch.zeros_like(k8), torch.zeros_like(v8)
    rows_k = k8[loc].view(torch.float8_e4m3fn).to(torch.bfloat16)
    ao.set_kv_buffer_fp8(kc, vc, loc, rows_k, rows_k)
    assert torch.equal(kc[loc], k8[loc])


def test_shared_prefix_segment_planner():
    from sglang_npu_b200.kernels import plan_shared_prefix_segments as plan

    # cfg 5 per GPU, head_dim 128: the two-query-tile prefix kernel takes 256 requests per CTA -> 8 CTAs per segment ->
    # 18 segments fill one wave of 148 SMs; head_dim 64 stays on 128-row CTAs: 16 CTAs per segment -> 9 segments
    seg = plan(30720, 256, 8, sm_count=148, sm_fraction=1.0)
    assert seg % 128 == 0 and (30720 + seg - 1) // seg == 18
    seg = plan(30720, 256, 8, sm_count=148, sm_fraction=1.0, head_dim=64)
    assert seg % 128 == 0 and (30720 + seg - 1) // seg == 9
    # never shorter than 512 tokens, always a multiple of 128, one segment when a segment already fills the chip
    assert plan(1024, 4, 2, sm_count=148, sm_fraction=1.0) == 512
    assert plan(5000, 64, 32 * 8, sm_count=148, sm_fraction=1.0) == 5120
    for p in (129, 1000, 4097, 30720):
        s = plan(p, 64, 32, sm_count=148, sm_fraction=1.0)
        assert s % 128 == 0 and s >= 512


def test_decode_split_planner_invariants():
    from sglang_npu_b200.srt.layers.attention.b200_backend import plan_decode_splits as plan

    for bs in (1, 7, 64, 256, 512):
        for hkv, group in ((8, 4), (1, 8), (2, 7), (4, 20)):
            for max_len in (1, 15, 300, 2048, 4096, 32768, 131072):
                for cap in (2048, 1024):
                    chunk, splits = plan(bs, hkv, group, max_len, 148, 0, cap)
                    assert chunk % 16 == 0 and 16 <= chunk <= cap
                    assert splits * chunk >= max_len and (splits - 1) * chunk < max_len
                    chunk, splits = plan(bs, hkv, group, max_len, 148, 0, cap, exact=False)
                    assert chunk % 128 == 0 and 128 <= chunk <= cap and splits * chunk >= max_len
    assert plan(64, 8, 4, 4096, 148) == (2048, 2)          # BASELINE config 2
    assert plan(64, 8, 4, 4096, 148, 0, 1024) == (1024, 4)  # ... with an FP8 KV cache
    assert plan(64, 8, 4, 4096, 148, forced_chunk=48) == (48, 86)
    # a captured graph planned under a bound a few decode steps past 4k keeps the 2048-token chunks of the 4k batch
    assert plan(64, 8, 4, 4096 + 32, 148, exact=False) == (2048, 3)
