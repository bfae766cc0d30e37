"""GPU parity: B200AttnBackend (through the C-ABI) vs the reference's golden vectors and the oracle."""
import os

import pytest
import torch

from oracle import attention_oracle as ao
from sglang_npu_b200 import kernels, synth
from tests import harness
from tests.util import assert_close_path, golden_files, load_golden

pytestmark = pytest.mark.gpu
DT = {"bf16": torch.bfloat16, "f16": torch.float16}


@pytest.mark.parametrize("path", golden_files("attn_"), ids=lambda p: p.split("attn_")[-1][:-3])
def test_backend_matches_reference_golden(cuda_device, path):
    """Same inputs the real reference (TorchNativeAttnBackend on CPU) was run on."""
    rec = load_golden(path)
    c = rec["case"]
    dtype = DT[c["dtype"]]
    runner, backend = harness.build_runner(cuda_device, rec["req_to_token"], rec["num_slots"], c["page"], c["hq"],
                                           c["hkv"], c["d"], dtype)
    pool = runner.token_to_kv_pool
    pool.k_buffer[0].copy_(rec["k_cache_before"])
    pool.v_buffer[0].copy_(rec["v_cache_before"])
    layer = harness.make_layer(c["hq"], c["hkv"], c["d"])
    if c["mode"] == "decode":
        fb = harness.decode_batch(runner, backend, rec["req_pool_indices"], rec["seq_lens"], rec["out_cache_loc"])
    else:
        fb = harness.extend_batch(runner, backend, rec["req_pool_indices"], rec["seq_lens"],
                                  rec["extend_prefix_lens"], rec["out_cache_loc"])
    q, k, v = (rec[n].to(cuda_device) for n in ("q", "k", "v"))
    o = layer(q, k, v, fb)
    torch.cuda.synchronize()
    # KV write: bit-exact
    assert torch.equal(pool.k_buffer[0].cpu().view(torch.int16), rec["k_cache_after"].view(torch.int16))
    assert torch.equal(pool.v_buffer[0].cpu().view(torch.int16), rec["v_cache_after"].view(torch.int16))
    assert o.shape == rec["o"].shape and o.dtype == dtype
    assert_close_path(o, rec["o"], f"golden {c['name']}")


def _decode_case(device, lens, hq, hkv, d, page, dtype, forced_chunk=0, seed=1, with_cpu_lens=True, logit_cap=0.0):
    pb = synth.make_paged_batch(lens, page, seed=seed)
    runner, backend = harness.build_runner(device, pb.req_to_token, pb.num_slots, page, hq, hkv, d, dtype,
                                           ctx=max(max(lens), 16))
    if forced_chunk:
        backend.forced_chunk = forced_chunk
    k0, v0 = synth.randn_kv(pb.num_slots, page, hkv, d, dtype, seed=seed + 10)
    pool = runner.token_to_kv_pool
    pool.k_buffer[0].copy_(k0)
    pool.v_buffer[0].copy_(v0)
    bs = len(lens)
    g = torch.Generator().manual_seed(seed + 20)
    q = torch.randn((bs, hq * d), generator=g).to(dtype)
    k = torch.randn((bs, hkv * d), generator=g).to(dtype)
    v = torch.randn((bs, hkv * d), generator=g).to(dtype)
    out_loc = torch.stack([pb.req_to_token[pb.req_pool_indices[i], max(lens[i] - 1, 0)] for i in range(bs)]).long()
    fb = harness.decode_batch(runner, backend, pb.req_pool_indices, pb.seq_lens, out_loc, with_cpu_lens)
    layer = harness.make_layer(hq, hkv, d, logit_cap=logit_cap)
    o = layer(q.to(device), k.to(device), v.to(device), fb)
    torch.cuda.synchronize()
    # oracle on CPU (fp32 evaluation of the reference formula on the post-write cache)
    ao.set_kv_buffer(k0, v0, out_loc, k.view(bs, hkv, d), v.view(bs, hkv, d))
    assert torch.equal(pool.k_buffer[0].cpu().view(torch.int16), k0.view(torch.int16))
    assert torch.equal(pool.v_buffer[0].cpu().view(torch.int16), v0.view(torch.int16))
    ref = ao.decode_attention(q.view(bs, hq, d), k0, v0, pb.req_to_token, pb.req_pool_indices, pb.seq_lens,
                              d ** -0.5, compute_dtype=torch.float32)
    return o.view(bs, hq, d), ref, backend


@pytest.mark.parametrize("hq,hkv,d", [(32, 8, 128), (14, 2, 64), (40, 8, 128), (8, 1, 128), (4, 4, 128),
                                      (64, 2, 64), (40, 2, 128)])
@pytest.mark.parametrize("page", [1, 16])
def test_decode_ragged_vs_oracle(cuda_device, hq, hkv, d, page):
    lens = [1, 2, 15, 16, 17, 255, 256, 257, 700, 1030]
    o, ref, _ = _decode_case(cuda_device, lens, hq, hkv, d, page, torch.bfloat16)
    assert_close_path(o, ref, f"decode hq{hq} hkv{hkv} d{d} page{page}")


@pytest.mark.parametrize("chunk", [16, 48, 256, 1024])
def test_decode_split_kv_chunks(cuda_device, chunk):
    """Multi-split path incl. chunk boundaries that cut pages and a last chunk of one token."""
    lens = [chunk + 1, 3 * chunk, 5 * chunk - 1, 7, 2 * chunk + 17]
    o, ref, backend = _decode_case(cuda_device, lens, 32, 8, 128, 16, torch.bfloat16, forced_chunk=chunk)
    assert backend.forward_metadata.max_splits >= 3
    assert_close_path(o, ref, f"decode split chunk {chunk}")


def test_decode_fp16_and_no_cpu_lens(cuda_device):
    o, ref, backend = _decode_case(cuda_device, [300, 1, 2049], 16, 4, 128, 16, torch.float16, with_cpu_lens=False)
    assert_close_path(o, ref, "decode fp16")


def test_decode_cfg1_shape(cuda_device):
    """BASELINE config 1: Qwen2.5-0.5B heads, bs=8, 512-token ctx, page_size=1."""
    o, ref, _ = _decode_case(cuda_device, [512] * 8, 14, 2, 64, 1, torch.bfloat16)
    assert_close_path(o, ref, "decode cfg1")


def test_decode_zero_length_request_and_workspace_error(cuda_device, lib):
    pb = synth.make_paged_batch([5, 40], 16, seed=3)
    dev = cuda_device
    k, v = synth.randn_kv(pb.num_slots, 16, 2, 128, torch.bfloat16)
    q = torch.randn(2, 4, 128, dtype=torch.bfloat16, device=dev)
    o = torch.full_like(q, 7.0)
    seq = torch.tensor([0, 40], dtype=torch.int64, device=dev)
    ws = torch.empty(kernels.decode_attn_workspace_bytes(2, 4, 128, 3), dtype=torch.uint8, device=dev)
    kernels.decode_attention_fwd(q, k.to(dev), v.to(dev), o, pb.req_to_token.to(dev), pb.req_pool_indices.to(dev), seq,
                                 ws, 16, 3, 128 ** -0.5)
    torch.cuda.synchronize()
    assert torch.all(o[0] == 0)
    with pytest.raises(RuntimeError, match="workspace"):
        kernels.decode_attention_fwd(q, k.to(dev), v.to(dev), o, pb.req_to_token.to(dev),
                                     pb.req_pool_indices.to(dev), seq, ws[:64], 16, 3, 128 ** -0.5)
    with pytest.raises(NotImplementedError):
        q96 = torch.randn(2, 4, 96, dtype=torch.bfloat16, device=dev)
        k96 = torch.randn(64, 2, 96, dtype=torch.bfloat16, device=dev)
        kernels.decode_attention_fwd(q96, k96, k96, torch.empty_like(q96), pb.req_to_token.to(dev),
                                     pb.req_pool_indices.to(dev), seq, ws, 16, 3, 1.0)


def _extend_case(device, lens, prefix, hq, hkv, d, page, dtype, seed=5, save_kv=True, causal=True):
    pb = synth.make_paged_batch(lens, page, seed=seed)
    runner, backend = harness.build_runner(device, pb.req_to_token, pb.num_slots, page, hq, hkv, d, dtype)
    k0, v0 = synth.randn_kv(pb.num_slots, page, hkv, d, dtype, seed=seed + 10)
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
    if not causal:
        from sglang_npu_b200.srt.layers.radix_attention import AttentionType
        layer.attn_type = AttentionType.ENCODER_ONLY
    o = layer(q.to(device), k.to(device), v.to(device), fb, save_kv_cache=save_kv)
    torch.cuda.synchronize()
    k1, v1 = k0.clone(), v0.clone()
    ao.set_kv_buffer(k1, v1, out_loc, k.view(T, hkv, d), v.view(T, hkv, d))
    if save_kv:
        assert torch.equal(pool.k_buffer[0].cpu().view(torch.int16), k1.view(torch.int16))
        assert torch.equal(pool.v_buffer[0].cpu().view(torch.int16), v1.view(torch.int16))
    else:  # cache untouched; new keys were read from the activations
        assert torch.equal(pool.k_buffer[0].cpu().view(torch.int16), k0.view(torch.int16))
    ref = ao.extend_attention(q.view(T, hq, d), k1, v1, pb.req_to_token, pb.req_pool_indices, pb.seq_lens,
                              prefix_t, ext, d ** -0.5, causal=causal, compute_dtype=torch.float32)
    return o.view(T, hq, d), ref


@pytest.mark.parametrize("hq,hkv,d", [(32, 8, 128), (14, 2, 64), (8, 8, 128)])
@pytest.mark.parametrize("page", [1, 16])
def test_extend_prefix_and_ragged(cuda_device, hq, hkv, d, page):
    lens = [1, 64, 65, 200, 333, 128]
    prefix = [0, 0, 64, 137, 16, 127]
    o, ref = _extend_case(cuda_device, lens, prefix, hq, hkv, d, page, torch.bfloat16)
    assert_close_path(o, ref, f"extend hq{hq} hkv{hkv} d{d} page{page}")


@pytest.mark.parametrize("page", [1, 16])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("causal", [True, False])
def test_extend_two_tile_tma_kernel(cuda_device, page, dtype, causal):
    """Requests with more than 128 new tokens take extend_attn_fa.cu (256 query rows per CTA, TMA page boxes):
    one- and two-tile blocks, ragged last KV tiles, prefixes that are not tile multiples, 1-row boxes (page 1)."""
    lens = [257, 206, 812, 1024, 130]
    prefix = [0, 77, 300, 24, 1] if causal else [0, 0, 0, 0, 0]
    o, ref = _extend_case(cuda_device, lens, prefix, 8, 2, 128, page, dtype, causal=causal)
    assert_close_path(o, ref, f"extend two-tile page{page} {dtype} causal={causal}")


def test_extend_two_tile_matches_one_tile_kernel_subprocess(cuda_device):
    """B200_EXTEND_FA=0 (read once per process) keeps the one-tile cp.async kernel: both kernels against the oracle on
    the same batch, and against each other."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import torch, sys\n"
        "from tests.test_attention_gpu import _extend_case\n"
        "o, ref = _extend_case(torch.device('cuda:0'), [700, 300], [100, 0], 32, 8, 128, 16, torch.bfloat16)\n"
        "torch.save(o.cpu(), sys.argv[1])\n"
    )
    outs = []
    for fa in ("1", "0"):
        path = f"/tmp/b200_extend_fa_{fa}_{os.getpid()}.pt"
        env = dict(os.environ, B200_EXTEND_FA=fa, PYTHONPATH=root)
        res = subprocess.run([sys.executable, "-c", code, path], capture_output=True, text=True, env=env, cwd=root, timeout=600)
        assert res.returncode == 0, res.stdout + res.stderr
        outs.append(torch.load(path))
        os.remove(path)
    assert_close_path(outs[0], outs[1].float(), "two-tile vs one-tile extend kernel")  # two bf16 roundings: 1 ulp at |o| ~ 2


def test_extend_without_saving_kv_reads_activations(cuda_device):
    o, ref = _extend_case(cuda_device, [100, 70], [30, 0], 8, 2, 128, 16, torch.bfloat16, save_kv=False)
    assert_close_path(o, ref, "extend save_kv_cache=False")


def test_extend_non_causal_encoder(cuda_device):
    o, ref = _extend_case(cuda_device, [90, 17], [0, 0], 8, 2, 128, 16, torch.float16, causal=False)
    assert_close_path(o, ref, "extend encoder-only")


def test_chunked_prefill_equals_single_prefill(cuda_device):
    """Config-3 behaviour: extending in 3 chunks (each chunk's prefix = what is cached so far) gives
    the same outputs as one extend of the whole prompt (size-independent property)."""
    dev = cuda_device
    hq, hkv, d, page, L = 8, 2, 128, 16, 400
    pb = synth.make_paged_batch([L], page, seed=11)
    g = torch.Generator().manual_seed(12)
    q = torch.randn((L, hq * d), generator=g).to(torch.bfloat16).to(dev)
    k = torch.randn((L, hkv * d), generator=g).to(torch.bfloat16).to(dev)
    v = torch.randn((L, hkv * d), generator=g).to(torch.bfloat16).to(dev)
    slots = pb.req_to_token[pb.req_pool_indices[0], :L].long()
    layer = harness.make_layer(hq, hkv, d)

    def run(bounds):
        runner, backend = harness.build_runner(dev, pb.req_to_token, pb.num_slots, page, hq, hkv, d, torch.bfloat16)
        outs = []
        for a, b in zip(bounds[:-1], bounds[1:]):
            fb = harness.extend_batch(runner, backend, pb.req_pool_indices, torch.tensor([b]), torch.tensor([a]),
                                      slots[a:b])
            outs.append(layer(q[a:b], k[a:b], v[a:b], fb))
        return torch.cat(outs)

    whole = run([0, L])
    chunks = run([0, 130, 257, L])
    torch.cuda.synchronize()
    assert_close_path(chunks, whole, "chunked vs whole prefill", max_abs=1e-2)


@pytest.mark.skipif(os.environ.get("B200_SKIP_FULLSIZE") == "1", reason="full-size case skipped by env")
def test_decode_full_size_properties(cuda_device):
    """BASELINE config 2 layer shape (bs=64, ctx=4096, Hq=32, Hkv=8, D=128, page=16), checked through
    size-independent properties: (a) V rows constant per (head, dim) => o equals that constant whatever
    the softmax weights / split merge do; (b) a shuffled page table that maps to the same logical KV
    gives bit-identical output (page gather is order-exact); (c) 64 sampled (request, head) pairs vs oracle."""
    dev = cuda_device
    bs, ctx, hq, hkv, d, page = 64, 4096, 32, 8, 128, 16
    pb = synth.make_paged_batch([ctx] * bs, page, seed=21)
    runner, backend = harness.build_runner(dev, pb.req_to_token, pb.num_slots, page, hq, hkv, d, torch.bfloat16)
    pool = runner.token_to_kv_pool
    kc, vc = synth.randn_kv_device(pb.num_slots, page, hkv, d, torch.bfloat16, dev, seed=22)
    pool.k_buffer[0].copy_(kc)
    const = torch.randn((hkv, d), device=dev).to(torch.bfloat16)
    pool.v_buffer[0].copy_(const.expand(pb.num_slots + page, hkv, d))
    q = torch.randn((bs, hq * d), device=dev).to(torch.bfloat16)
    layer = harness.make_layer(hq, hkv, d)
    out_loc = torch.zeros(bs, dtype=torch.int64)
    fb = harness.decode_batch(runner, backend, pb.req_pool_indices, pb.seq_lens, out_loc)
    assert backend.forward_metadata.max_splits > 1
    o = layer(q, None, None, fb, save_kv_cache=False).view(bs, hq, d)
    expect = const.float().repeat_interleave(hq // hkv, dim=0).unsqueeze(0).expand(bs, hq, d)
    assert (o.float() - expect).abs().max().item() <= 2e-2
    # (c) real V, sampled pairs vs oracle
    pool.v_buffer[0].copy_(vc)
    o2 = layer(q, None, None, fb, save_kv_cache=False).view(bs, hq, d)
    torch.cuda.synchronize()
    sel = [0, 17, 63]
    kcpu, vcpu = kc.cpu(), vc.cpu()
    ref = ao.decode_attention(q.view(bs, hq, d)[sel].cpu(), kcpu, vcpu, pb.req_to_token, pb.req_pool_indices[sel],
                              pb.seq_lens[sel], d ** -0.5, compute_dtype=torch.float32)
    assert_close_path(o2[sel], ref, "full-size sampled requests")
    # (b) idempotence: same launch twice is bit-identical (no atomics / order dependence)
    o3 = layer(q, None, None, fb, save_kv_cache=False).view(bs, hq, d)
    assert torch.equal(o2, o3)


@pytest.mark.parametrize("hq,hkv,d", [(32, 8, 128), (14, 2, 64), (40, 8, 128)])
@pytest.mark.parametrize("forced_chunk", [0, 64])
def test_decode_int8_out_equals_unfused_quant(cuda_device, hq, hkv, d, forced_chunk):
    """Stage 2 fused with per_token_quant_int8 (b200_decode_attn_int8): same bits as the two kernels back to back,
    for requests merged from several chunks and for requests stage 1 finalised itself (one chunk)."""
    dev = cuda_device
    lens = [1, 2, 15, 16, 17, 255, 256, 257, 700, 1030]
    page, dtype = 16, torch.bfloat16
    pb = synth.make_paged_batch(lens, page, seed=4)
    runner, backend = harness.build_runner(dev, pb.req_to_token, pb.num_slots, page, hq, hkv, d, dtype, ctx=max(lens))
    if forced_chunk:
        backend.forced_chunk = forced_chunk
    k0, v0 = synth.randn_kv(pb.num_slots, page, hkv, d, dtype, seed=14)
    runner.token_to_kv_pool.k_buffer[0].copy_(k0)
    runner.token_to_kv_pool.v_buffer[0].copy_(v0)
    bs = len(lens)
    q = torch.randn((bs, hq * d), generator=torch.Generator().manual_seed(5)).to(dtype).to(dev)
    out_loc = torch.stack([pb.req_to_token[pb.req_pool_indices[i], lens[i] - 1] for i in range(bs)]).long()
    fb = harness.decode_batch(runner, backend, pb.req_pool_indices, pb.seq_lens, out_loc)
    layer = harness.make_layer(hq, hkv, d)
    o = layer(q, None, None, fb, save_kv_cache=False)
    act = layer(q, None, None, fb, save_kv_cache=False, int8_out=True)
    xq, xs = kernels.per_token_quant_int8(o)
    torch.cuda.synchronize()
    assert isinstance(act, kernels.Int8Activation)
    assert torch.equal(act.x_q.cpu(), xq.cpu())
    assert torch.equal(act.x_scale.cpu().view(torch.int32), xs.cpu().view(torch.int32))


def _capped_attention_rows(q, k_all, v_all, positions, hq, hkv, d, cap):
    """softmax(cap * tanh(q k^T * scale / cap)) v in fp32 for single query rows; row i sees keys [0, positions[i]].
    The logit-cap formula of the reference's Triton kernels (triton_ops/extend_attention.py:164-165, :242-243;
    decode_attention.py:122-123) — torch_native ignores logit_cap, so this restatement is the yardstick."""
    grp = hq // hkv
    out = torch.empty(q.shape[0], hq, d)
    for i, p in enumerate(positions):
        k = k_all[: p + 1].float()
        v = v_all[: p + 1].float()
        qi = q[i].float().view(hkv, grp, d)
        s = torch.einsum("hgd,thd->hgt", qi, k) * d ** -0.5
        s = cap * torch.tanh(s / cap)
        out[i] = torch.einsum("hgt,thd->hgd", torch.softmax(s, dim=-1), v).reshape(hq, d)
    return out


@pytest.mark.parametrize("new,prefix", [(300, 77), (100, 0)])
def test_logit_cap_extend_and_decode(cuda_device, new, prefix):
    """logit_cap > 0 (Gemma-style soft capping): the two-tile TMA extend kernel (300 new tokens), the one-tile kernel
    (100) and the decode kernel apply cap * tanh(s / cap) before the softmax."""
    dev = cuda_device
    hq, hkv, d, page, cap = 8, 2, 128, 16, 5.0
    total = prefix + new
    pb = synth.make_paged_batch([total + 1], page, seed=41)
    runner, backend = harness.build_runner(dev, pb.req_to_token, pb.num_slots, page, hq, hkv, d, torch.bfloat16, ctx=total + 1)
    pool = runner.token_to_kv_pool
    kc, vc = synth.randn_kv_device(pb.num_slots, page, hkv, d, torch.bfloat16, dev, seed=42)
    pool.k_buffer[0].copy_(kc * 2)  # larger scores so the cap bites
    pool.v_buffer[0].copy_(vc)
    g = torch.Generator().manual_seed(43)
    q = (torch.randn(new, hq * d, generator=g) * 2).to(torch.bfloat16).to(dev)
    k = (torch.randn(new, hkv * d, generator=g) * 2).to(torch.bfloat16).to(dev)
    v = torch.randn(new, hkv * d, generator=g).to(torch.bfloat16).to(dev)
    slots = pb.req_to_token[pb.req_pool_indices[0], : total + 1].long()
    fb = harness.extend_batch(runner, backend, pb.req_pool_indices, torch.tensor([total]), torch.tensor([prefix]),
                              slots[prefix:total])
    layer = harness.make_layer(hq, hkv, d, logit_cap=cap)
    o = layer(q, k, v, fb).view(new, hq, d)
    torch.cuda.synchronize()
    k_all = pool.k_buffer[0][slots.to(dev)].cpu()
    v_all = pool.v_buffer[0][slots.to(dev)].cpu()
    rows = sorted(set([0, 1, new // 2, new - 2, new - 1] + list(range(120, min(new, 140)))))
    rows = [r for r in rows if r < new]
    ref = _capped_attention_rows(q.cpu().view(new, hq, d)[rows], k_all, v_all, [prefix + r for r in rows], hq, hkv, d, cap)
    assert_close_path(o[rows], ref, f"extend with logit_cap, {new} new tokens")
    uncapped = _attn_rows_plain(q.cpu().view(new, hq, d)[rows], k_all, v_all, [prefix + r for r in rows], hq, hkv, d)
    assert (uncapped - ref).abs().max().item() > 5e-2, "the cap must change the result for this test to mean anything"
    # one decode step on top
    qd = (torch.randn(1, hq * d, generator=g) * 2).to(torch.bfloat16).to(dev)
    kd = (torch.randn(1, hkv * d, generator=g) * 2).to(torch.bfloat16).to(dev)
    vd = torch.randn(1, hkv * d, generator=g).to(torch.bfloat16).to(dev)
    fbd = harness.decode_batch(runner, backend, pb.req_pool_indices, torch.tensor([total + 1]), slots[total: total + 1])
    od = layer(qd, kd, vd, fbd).view(1, hq, d)
    torch.cuda.synchronize()
    k_all = pool.k_buffer[0][slots.to(dev)].cpu()
    v_all = pool.v_buffer[0][slots.to(dev)].cpu()
    refd = _capped_attention_rows(qd.cpu().view(1, hq, d), k_all, v_all, [total], hq, hkv, d, cap)
    assert_close_path(od, refd, "decode with logit_cap")


def _attn_rows_plain(q, k_all, v_all, positions, hq, hkv, d):
    grp = hq // hkv
    out = torch.empty(q.shape[0], hq, d)
    for i, p in enumerate(positions):
        s = torch.einsum("hgd,thd->hgt", q[i].float().view(hkv, grp, d), k_all[: p + 1].float()) * d ** -0.5
        out[i] = torch.einsum("hgt,thd->hgd", torch.softmax(s, dim=-1), v_all[: p + 1].float()).reshape(hq, d)
    return out
