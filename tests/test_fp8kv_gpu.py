"""GPU parity for the FP8 (e4m3) KV cache: the cast-and-scatter KV write is bit-exact against the
reference's MHATokenToKVPool golden vectors and the oracle; decode attention over the fp8 cache agrees with the oracle
within the north_star tolerance."""
import pytest
import torch

from oracle import attention_oracle as ao
from sglang_npu_b200 import kernels, synth
from tests import harness
from tests.util import assert_close_path, golden_files, load_golden

pytestmark = pytest.mark.gpu
FP8 = torch.float8_e4m3fn


def test_fp8_kv_write_matches_reference_golden(cuda_device):
    for rec in load_golden(golden_files("kvwrite_fp8")[0]):
        pool = harness.MHATokenToKVPool(rec["slots"], rec["page"], FP8, rec["hkv"], rec["d"], 1, cuda_device)
        assert pool.k_buffer[0].dtype == torch.uint8
        layer = harness.make_layer(rec["hkv"], rec["hkv"], rec["d"])
        k, v = rec["k"].to(cuda_device), rec["v"].to(cuda_device)
        pool.set_kv_buffer(layer, rec["loc"].to(cuda_device), k, v, rec["k_scale"], rec["v_scale"])
        torch.cuda.synchronize()
        assert torch.equal(pool.k_buffer[0].cpu(), rec["k_cache_after"])
        assert torch.equal(pool.v_buffer[0].cpu(), rec["v_cache_after"])
        assert torch.equal(k.cpu().view(torch.int16), rec["k"].view(torch.int16))  # caller's rows left untouched


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("scales", [(None, None), (0.43, 1.9)])
def test_fp8_kv_write_random_vs_oracle(cuda_device, dtype, scales):
    """Every bf16/f16 bit pattern class at scale: 4096 rows x 1024 values, int32 loc, strided source rows."""
    g = torch.Generator().manual_seed(5)
    T, hkv, d, slots = 4096, 8, 128, 8192
    big = (torch.randn((T, 3 * hkv * d), generator=g) * 60).to(dtype)
    k, v = big[:, : hkv * d], big[:, hkv * d: 2 * hkv * d]  # views with a 3x token stride, like a fused qkv output
    loc = (torch.randperm(slots, generator=g)[:T] + 16).to(torch.int32)
    kc = torch.zeros((slots + 16, hkv, d), dtype=torch.uint8)
    vc = torch.zeros_like(kc)
    ao.set_kv_buffer_fp8(kc, vc, loc.long(), k.reshape(T, hkv, d), v.reshape(T, hkv, d), *scales)
    kd = torch.zeros_like(kc, device=cuda_device)
    vd = torch.zeros_like(kd)
    bigd = big.to(cuda_device)
    kernels.set_kv_buffer(kd, vd, loc.to(cuda_device), bigd[:, : hkv * d], bigd[:, hkv * d: 2 * hkv * d], *scales)
    torch.cuda.synchronize()
    assert torch.equal(kd.cpu(), kc) and torch.equal(vd.cpu(), vc)


def _decode_case(device, lens, hq, hkv, d, page, dtype, scales=(None, None), forced_chunk=0, seed=1, int8_out=False):
    pb = synth.make_paged_batch(lens, page, seed=seed)
    runner, backend = harness.build_runner(device, pb.req_to_token, pb.num_slots, page, hq, hkv, d, FP8,
                                           ctx=max(max(lens), 16))
    if forced_chunk:
        backend.forced_chunk = forced_chunk
    pool = runner.token_to_kv_pool
    k0, v0 = synth.randn_kv(pb.num_slots, page, hkv, d, dtype, seed=seed + 10)
    k0 = k0.to(FP8).view(torch.uint8)
    v0 = v0.to(FP8).view(torch.uint8)
    pool.k_buffer[0].copy_(k0)
    pool.v_buffer[0].copy_(v0)
    bs = len(lens)
    g = torch.Generator().manual_seed(seed + 20)
    q = torch.randn((bs, hq * d), generator=g).to(dtype)
    k = torch.randn((bs, hkv * d), generator=g).to(dtype)
    v = torch.randn((bs, hkv * d), generator=g).to(dtype)
    out_loc = torch.stack([pb.req_to_token[pb.req_pool_indices[i], max(lens[i] - 1, 0)] for i in range(bs)]).long()
    fb = harness.decode_batch(runner, backend, pb.req_pool_indices, pb.seq_lens, out_loc)
    layer = harness.make_layer(hq, hkv, d)
    layer.k_scale_float, layer.v_scale_float = scales
    if scales[0] is not None:
        layer.k_scale = torch.tensor(scales[0])
        layer.v_scale = torch.tensor(scales[1])
    kw = dict(int8_out=True) if int8_out else {}
    o = layer(q.to(device), k.to(device), v.to(device), fb, **kw)
    torch.cuda.synchronize()
    ao.set_kv_buffer_fp8(k0, v0, out_loc, k.view(bs, hkv, d), v.view(bs, hkv, d), *scales)
    assert torch.equal(pool.k_buffer[0].cpu(), k0) and torch.equal(pool.v_buffer[0].cpu(), v0)
    ref = ao.decode_attention_fp8kv(q.view(bs, hq, d), k0, v0, pb.req_to_token, pb.req_pool_indices, pb.seq_lens,
                                    d ** -0.5, *scales)
    return o, ref, backend


@pytest.mark.parametrize("hq,hkv,d", [(32, 8, 128), (14, 2, 64), (40, 8, 128), (8, 1, 128), (4, 4, 128), (64, 2, 64),
                                      (40, 2, 128)])
@pytest.mark.parametrize("page", [1, 16])
def test_fp8kv_decode_ragged_vs_oracle(cuda_device, hq, hkv, d, page):
    lens = [1, 2, 15, 16, 17, 255, 256, 257, 700, 1030]
    o, ref, _ = _decode_case(cuda_device, lens, hq, hkv, d, page, torch.bfloat16)
    assert_close_path(o.view(ref.shape), ref, f"fp8kv decode hq{hq} hkv{hkv} d{d} page{page}")


@pytest.mark.parametrize("chunk", [16, 48, 256])
def test_fp8kv_decode_split_chunks_f16_and_scales(cuda_device, chunk):
    lens = [chunk + 1, 3 * chunk, 5 * chunk - 1, 7, 2 * chunk + 17]
    o, ref, backend = _decode_case(cuda_device, lens, 32, 8, 128, 16, torch.float16, scales=(0.5, 1.75),
                                   forced_chunk=chunk)
    assert backend.forward_metadata.max_splits >= 3
    assert_close_path(o.view(ref.shape), ref, f"fp8kv decode f16 scales chunk {chunk}")


def test_fp8kv_decode_int8_out_equals_unfused_quant(cuda_device):
    lens = [33, 700, 1, 256]
    act, ref, _ = _decode_case(cuda_device, lens, 32, 8, 128, 16, torch.bfloat16, int8_out=True)
    o, _, _ = _decode_case(cuda_device, lens, 32, 8, 128, 16, torch.bfloat16)
    xq, xs = kernels.per_token_quant_int8(o)
    assert torch.equal(act[0], xq) and torch.equal(act[1], xs)


def test_fp8kv_decode_cfg2_layer_properties(cuda_device):
    """BASELINE config 2 shapes (bs 64, ctx 4096, page 16) over an fp8 cache: permutation invariance of the page
    layout and split-count invariance at full size; a slice of requests against the oracle."""
    lens = [4096] * 64
    o_a, _, _ = _decode_case_fast(cuda_device, lens, seed=3, forced_chunk=0)
    o_b, _, _ = _decode_case_fast(cuda_device, lens, seed=3, forced_chunk=512)
    assert_close_path(o_a, o_b, "fp8kv cfg2 split invariance", max_abs=2e-3)


def _decode_case_fast(device, lens, seed, forced_chunk):
    hq, hkv, d, page = 32, 8, 128, 16
    pb = synth.make_paged_batch(lens, page, seed=seed)
    runner, backend = harness.build_runner(device, pb.req_to_token, pb.num_slots, page, hq, hkv, d, FP8,
                                           ctx=max(lens))
    if forced_chunk:
        backend.forced_chunk = forced_chunk
    pool = runner.token_to_kv_pool
    g = torch.Generator(device=device).manual_seed(seed)
    for buf in (pool.k_buffer[0], pool.v_buffer[0]):
        buf.copy_(torch.randn(buf.shape, generator=g, device=device, dtype=torch.bfloat16).to(FP8).view(torch.uint8))
    bs = len(lens)
    q = torch.randn((bs, hq * d), generator=g, device=device, dtype=torch.bfloat16)
    out_loc = torch.stack([pb.req_to_token[pb.req_pool_indices[i], lens[i] - 1] for i in range(bs)]).long()
    fb = harness.decode_batch(runner, backend, pb.req_pool_indices, pb.seq_lens, out_loc)
    layer = harness.make_layer(hq, hkv, d)
    o = layer(q, None, None, fb, save_kv_cache=False)
    torch.cuda.synchronize()
    # three requests against the CPU oracle
    sel = [0, 31, 63]
    ref = ao.decode_attention_fp8kv(q.cpu().view(bs, hq, d)[sel], pool.k_buffer[0].cpu(), pool.v_buffer[0].cpu(),
                                    pb.req_to_token, pb.req_pool_indices[sel], pb.seq_lens[sel], d ** -0.5)
    assert_close_path(o.view(bs, hq, d)[sel], ref, "fp8kv cfg2 rows vs oracle")
    return o, ref, backend


def _extend_case(device, lens, prefix, hq, hkv, d, page, dtype, scales=(None, None), seed=5, causal=True):
    pb = synth.make_paged_batch(lens, page, seed=seed)
    runner, backend = harness.build_runner(device, pb.req_to_token, pb.num_slots, page, hq, hkv, d, FP8)
    k0, v0 = synth.randn_kv(pb.num_slots, page, hkv, d, dtype, seed=seed + 10)
    k0 = k0.to(FP8).view(torch.uint8)
    v0 = v0.to(FP8).view(torch.uint8)
    pool = runner.token_to_kv_pool
    pool.k_buffer[0].copy_(k0)
    pool.v_buffer[0].copy_(v0)
    bs = len(lens)
    prefix_t = torch.tensor(prefix, dtype=torch.int32)
    ext = pb.seq_lens.to(torch.int32) - prefix_t
    T = int(ext.sum())
    g = torch.Generator().manual_seed(seed + 20)
    q = torch.randn((T, hq * d), generator=g).to(dtype)
    k = torch.randn((T, hkv * d), generator=g).to(dtype)
    v = torch.randn((T, hkv * d), generator=g).to(dtype)
    out_loc = torch.cat([pb.req_to_token[pb.req_pool_indices[i], prefix[i]: lens[i]] for i in range(bs)]).long()
    fb = harness.extend_batch(runner, backend, pb.req_pool_indices, pb.seq_lens, prefix_t, out_loc)
    layer = harness.make_layer(hq, hkv, d)
    layer.k_scale_float, layer.v_scale_float = scales
    if scales[0] is not None:
        layer.k_scale, layer.v_scale = torch.tensor(scales[0]), torch.tensor(scales[1])
    if not causal:
        from sglang_npu_b200.srt.layers.radix_attention import AttentionType
        layer.attn_type = AttentionType.ENCODER_ONLY
    o = layer(q.to(device), k.to(device), v.to(device), fb)
    torch.cuda.synchronize()
    ref = ao.extend_attention_fp8kv(q.view(T, hq, d), k.view(T, hkv, d), v.view(T, hkv, d), k0, v0, pb.req_to_token,
                                    pb.req_pool_indices, pb.seq_lens, prefix_t, ext, out_loc, d ** -0.5, *scales,
                                    causal=causal)
    # the step's KV rows were written quantised, bit-exact
    ao.set_kv_buffer_fp8(k0, v0, out_loc, k.view(T, hkv, d), v.view(T, hkv, d), *scales)
    assert torch.equal(pool.k_buffer[0].cpu(), k0) and torch.equal(pool.v_buffer[0].cpu(), v0)
    return o.view(T, hq, d), ref


@pytest.mark.parametrize("hq,hkv,d", [(32, 8, 128), (14, 2, 64)])
@pytest.mark.parametrize("page", [1, 16])
def test_fp8kv_extend_prefix_and_ragged(cuda_device, hq, hkv, d, page):
    lens = [70, 33, 16, 300, 129, 1]
    prefix = [32, 0, 15, 128, 1, 0]
    o, ref = _extend_case(cuda_device, lens, prefix, hq, hkv, d, page, torch.bfloat16)
    assert_close_path(o, ref, f"fp8kv extend hq{hq} hkv{hkv} d{d} page{page}")


def test_fp8kv_extend_scales_f16_and_no_prefix(cuda_device):
    o, ref = _extend_case(cuda_device, [200, 257], [64, 129], 16, 4, 128, 16, torch.float16, scales=(0.5, 2.0))
    assert_close_path(o, ref, "fp8kv extend f16 scales")
    o, ref = _extend_case(cuda_device, [40, 9], [0, 0], 8, 2, 128, 16, torch.bfloat16)
    assert_close_path(o, ref, "fp8kv extend, no cached prefix at all")


def test_fp8kv_chunked_prefill_cfg3_like(cuda_device):
    """A 2048-token chunk on top of a 1536-token cached prefix (config 3's chunked prefill with prefix reuse, scaled
    down), fp8 cache."""
    o, ref = _extend_case(cuda_device, [3584], [1536], 32, 8, 128, 16, torch.bfloat16)
    assert_close_path(o, ref, "fp8kv chunked prefill")


def test_fp8_kv_gather_dequant_exact(cuda_device):
    g = torch.Generator().manual_seed(3)
    hkv, d, page = 2, 128, 16
    lens = [50, 17, 0, 33]
    pb = synth.make_paged_batch([max(n, 1) for n in lens], page, seed=9)
    cache = torch.randint(0, 256, (pb.num_slots + page, hkv, d), generator=g, dtype=torch.int32).to(torch.uint8)
    cache[(cache & 0x7F) == 0x7F] = 0x3C  # no NaN codes
    vcache = cache.flip(0).contiguous()
    prefix = torch.tensor(lens, dtype=torch.int32)
    offs = (torch.cumsum(prefix, 0) - prefix).to(torch.int32)
    dev = cuda_device
    ko = torch.zeros((int(prefix.sum()), hkv, d), dtype=torch.bfloat16, device=dev)
    vo = torch.zeros_like(ko)
    kernels.fp8_kv_gather_dequant(ko, vo, cache.to(dev), vcache.to(dev), pb.req_to_token.to(dev),
                                  pb.req_pool_indices.to(dev), prefix.to(dev), offs.to(dev), max(lens))
    torch.cuda.synchronize()
    for b, n in enumerate(lens):
        toks = pb.req_to_token[pb.req_pool_indices[b], :n].long()
        exp_k = cache[toks].view(FP8).to(torch.bfloat16)
        exp_v = vcache[toks].view(FP8).to(torch.bfloat16)
        assert torch.equal(ko[int(offs[b]): int(offs[b]) + n].cpu().view(torch.int16), exp_k.view(torch.int16))
        assert torch.equal(vo[int(offs[b]): int(offs[b]) + n].cpu().view(torch.int16), exp_v.view(torch.int16))
