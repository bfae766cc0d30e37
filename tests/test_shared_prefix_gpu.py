"""GPU parity for the shared-prefix (cascade) decode (BASELINE config 5): prefix pass on tcgen05 +
per-request split-KV decode + one LSE merge must equal plain decode attention — checked against the oracle
(TorchNativeAttnBackend's formula, which knows nothing about shared prefixes) and against the plain decode kernel."""
import pytest
import torch

from oracle import attention_oracle as ao
from sglang_npu_b200 import kernels, synth
from tests import harness
from tests.util import assert_close_path

pytestmark = pytest.mark.gpu


def _case(device, lens, shared, hq, hkv, d, page, dtype=torch.bfloat16, seed=1, int8_out=False, oracle_rows=None,
          forced_chunk=0):
    pb = synth.make_paged_batch(lens, page, seed=seed, shared_prefix=shared // page * page)
    outs = []
    bs = len(lens)
    g = torch.Generator().manual_seed(seed + 20)
    q = torch.randn((bs, hq * d), generator=g).to(dtype)
    k0 = v0 = None
    for use_shared in (True, False):
        runner, backend = harness.build_runner(device, pb.req_to_token, pb.num_slots, page, hq, hkv, d, dtype,
                                               ctx=max(max(lens), 16))
        backend.min_shared_prefix = 128
        backend.forced_chunk = forced_chunk
        pool = runner.token_to_kv_pool
        if k0 is None:
            k0, v0 = synth.randn_kv_device(pb.num_slots, page, hkv, d, dtype, device, seed=seed + 10)
        pool.k_buffer[0].copy_(k0)
        pool.v_buffer[0].copy_(v0)
        out_loc = torch.stack([pb.req_to_token[pb.req_pool_indices[i], lens[i] - 1] for i in range(bs)]).long()
        fb = harness.decode_batch(runner, backend, pb.req_pool_indices, pb.seq_lens, out_loc,
                                  shared_prefix_len=shared if use_shared else 0)
        assert (backend.forward_metadata.shared_prefix_len > 0) == use_shared
        layer = harness.make_layer(hq, hkv, d)
        kw = dict(int8_out=True) if int8_out else {}
        o = layer(q.to(device), None, None, fb, save_kv_cache=False, **kw)
        torch.cuda.synchronize()
        outs.append(o)
    if int8_out:
        return outs
    rows = list(range(bs)) if oracle_rows is None else oracle_rows
    ref = ao.decode_attention(q.view(bs, hq, d)[rows], k0.cpu(), v0.cpu(), pb.req_to_token, pb.req_pool_indices[rows],
                              pb.seq_lens[rows], d ** -0.5, compute_dtype=torch.float32)
    assert_close_path(outs[0].view(bs, hq, d)[rows], ref, "shared-prefix decode vs oracle")
    # same math, different association: agree far inside the north_star tolerance
    assert_close_path(outs[0], outs[1], "shared-prefix vs plain decode kernel", max_abs=4e-3)
    return outs


@pytest.mark.parametrize("hq,hkv,d", [(32, 8, 128), (14, 2, 64), (8, 1, 128), (40, 8, 128)])
@pytest.mark.parametrize("page", [1, 16])
def test_shared_prefix_ragged_vs_oracle(cuda_device, hq, hkv, d, page):
    lens = [257, 258, 300, 1030, 700, 511, 512, 513]
    _case(cuda_device, lens, 256, hq, hkv, d, page)


def test_shared_prefix_unaligned_prefix_and_fp16(cuda_device):
    """A 700-token common prefix: 640 go through the prefix pass, 60 stay per-request."""
    _case(cuda_device, [701, 900, 1500], 700, 16, 4, 128, 16, dtype=torch.float16)


def test_shared_prefix_many_segments_and_two_row_blocks(cuda_device):
    """130 requests (two 128-row blocks in the prefix pass), 4096 shared tokens cut into several segments."""
    lens = [4096 + 1 + (i * 37) % 300 for i in range(130)]
    _case(cuda_device, lens, 4096, 8, 2, 128, 16, oracle_rows=[0, 1, 64, 127, 128, 129])


def test_shared_prefix_multi_split_unique_part(cuda_device):
    _case(cuda_device, [1024 + 600, 1024 + 1, 1024 + 333], 1024, 32, 8, 128, 16, forced_chunk=64)


def test_shared_prefix_int8_out_equals_unfused_quant(cuda_device):
    lens = [600, 700, 1025, 513]
    act, _ = _case(cuda_device, lens, 512, 32, 8, 128, 16, int8_out=True)
    o, _ = _case(cuda_device, lens, 512, 32, 8, 128, 16)
    xq, xs = kernels.per_token_quant_int8(o)
    assert torch.equal(act[0], xq) and torch.equal(act[1], xs)


def test_shared_prefix_cfg5_per_gpu_shape(cuda_device):
    """BASELINE config 5 per GPU at TP=8: Hq 8, Hkv 1, D 128, bs 256, 30,720 shared + up to 2,048 own tokens."""
    g = torch.Generator().manual_seed(0)
    lens = (30720 + torch.randint(1, 2049, (256,), generator=g)).tolist()
    _case(cuda_device, lens, 30720, 8, 1, 128, 16, oracle_rows=[0, 100, 255])


def test_short_or_absent_prefix_takes_plain_path(cuda_device):
    pb = synth.make_paged_batch([300, 400], 16, seed=3, shared_prefix=64)
    runner, backend = harness.build_runner(cuda_device, pb.req_to_token, pb.num_slots, 16, 8, 2, 128, torch.bfloat16)
    out_loc = torch.zeros(2, dtype=torch.int64)
    harness.decode_batch(runner, backend, pb.req_pool_indices, pb.seq_lens, out_loc, shared_prefix_len=64)
    assert backend.forward_metadata.shared_prefix_len == 0  # below B200_MIN_SHARED_PREFIX


def test_shared_prefix_decode_is_cuda_graph_capturable(cuda_device):
    """The fork/join of the prefix pass onto the side stream is plain event record/wait: capture + replay must
    reproduce the eager result bit for bit."""
    dev = cuda_device
    hq, hkv, d, page, shared = 8, 2, 128, 16, 512
    lens = [600, 700, 513, 1025]
    bs = len(lens)
    pb = synth.make_paged_batch(lens, page, seed=7, shared_prefix=shared)
    k, v = synth.randn_kv_device(pb.num_slots, page, hkv, d, torch.bfloat16, dev, seed=8)
    q = torch.randn(bs, hq, d, device=dev).to(torch.bfloat16)
    r2t, rpi, sl = pb.req_to_token.to(dev), pb.req_pool_indices.to(dev), pb.seq_lens.to(dev)
    seg = kernels.plan_shared_prefix_segments(shared, bs, hq)
    nseg = (shared + seg - 1) // seg
    chunk, splits = 256, 3
    ws = torch.empty(kernels.decode_attn_workspace_bytes(bs, hq, d, nseg + splits), dtype=torch.uint8, device=dev)
    o_eager, o_graph = torch.empty_like(q), torch.empty_like(q)

    def run(o):
        kernels.decode_attention_shared_prefix_fwd(q, k, v, o, r2t, rpi, sl, shared, seg, ws, chunk, splits, d ** -0.5)

    run(o_eager)
    torch.cuda.synchronize()
    side = torch.cuda.Stream(device=dev)
    side.wait_stream(torch.cuda.current_stream(dev))
    with torch.cuda.stream(side):
        run(o_graph)
    torch.cuda.current_stream(dev).wait_stream(side)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        run(o_graph)
    o_graph.zero_()
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    assert torch.equal(o_graph.view(torch.int16), o_eager.view(torch.int16))


def _fp8_case(device, lens, shared, hq, hkv, d, page, dtype=torch.bfloat16, scales=(None, None), seed=3, graph=False):
    """Shared-prefix decode over an e4m3 KV cache (VERDICT r1 missing #7): the staged prefix pass + the FP8-KV decode
    kernel vs (a) the FP8-KV decode oracle, (b) the plain FP8-KV decode kernel on the same batch."""
    FP8 = torch.float8_e4m3fn
    pb = synth.make_paged_batch(lens, page, seed=seed, shared_prefix=shared // page * page)
    bs = len(lens)
    g = torch.Generator().manual_seed(seed + 20)
    q = torch.randn((bs, hq * d), generator=g).to(dtype)
    k0, v0 = synth.randn_kv(pb.num_slots, page, hkv, d, dtype, seed=seed + 10)
    k0, v0 = k0.to(FP8).view(torch.uint8), v0.to(FP8).view(torch.uint8)
    outs = []
    for use_shared in (True, False):
        runner, backend = harness.build_runner(device, pb.req_to_token, pb.num_slots, page, hq, hkv, d, FP8,
                                               ctx=max(max(lens), 16))
        backend.min_shared_prefix = 128
        pool = runner.token_to_kv_pool
        pool.k_buffer[0].copy_(k0)
        pool.v_buffer[0].copy_(v0)
        out_loc = torch.stack([pb.req_to_token[pb.req_pool_indices[i], lens[i] - 1] for i in range(bs)]).long()
        fb = harness.decode_batch(runner, backend, pb.req_pool_indices, pb.seq_lens, out_loc,
                                  shared_prefix_len=shared if use_shared else 0)
        assert (backend.forward_metadata.shared_prefix_len > 0) == use_shared
        layer = harness.make_layer(hq, hkv, d)
        layer.k_scale_float, layer.v_scale_float = scales
        if scales[0] is not None:
            layer.k_scale, layer.v_scale = torch.tensor(scales[0]), torch.tensor(scales[1])
        qd = q.to(device)
        o = layer(qd, None, None, fb, save_kv_cache=False)
        if graph and use_shared:  # the staging gather + fork / join under capture
            torch.cuda.synchronize()
            side = torch.cuda.Stream(device=device)
            side.wait_stream(torch.cuda.current_stream(device))
            with torch.cuda.stream(side):
                layer(qd, None, None, fb, save_kv_cache=False)
            torch.cuda.current_stream(device).wait_stream(side)
            torch.cuda.synchronize()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                og = layer(qd, None, None, fb, save_kv_cache=False)
            og.zero_()
            gr.replay()
            torch.cuda.synchronize()
            assert torch.equal(og.view(torch.int16), o.view(torch.int16))
        torch.cuda.synchronize()
        outs.append(o)
    ref = ao.decode_attention_fp8kv(q.view(bs, hq, d), k0, v0, pb.req_to_token, pb.req_pool_indices, pb.seq_lens,
                                    d ** -0.5, *scales)
    assert_close_path(outs[0].view(ref.shape), ref, "fp8-KV shared-prefix decode vs oracle")
    assert_close_path(outs[0], outs[1], "fp8-KV shared-prefix vs plain fp8-KV decode kernel", max_abs=8e-3)


@pytest.mark.parametrize("hq,hkv,d", [(32, 8, 128), (8, 1, 128), (14, 2, 64)])
@pytest.mark.parametrize("page", [1, 16])
def test_shared_prefix_fp8_kv_ragged_vs_oracle(cuda_device, hq, hkv, d, page):
    _fp8_case(cuda_device, [257, 258, 300, 1030, 700, 511, 512, 513], 256, hq, hkv, d, page)


def test_shared_prefix_fp8_kv_scales_f16_unaligned_prefix(cuda_device):
    _fp8_case(cuda_device, [701, 900, 1500], 700, 16, 4, 128, 16, dtype=torch.float16, scales=(0.5, 1.75))


def test_shared_prefix_fp8_kv_cuda_graph(cuda_device):
    _fp8_case(cuda_device, [600, 700, 513, 1025], 512, 8, 2, 128, 16, graph=True)


def test_shared_prefix_fp8_kv_cfg5_per_gpu_shape(cuda_device):
    g = torch.Generator().manual_seed(0)
    lens = (30720 + torch.randint(1, 2049, (64,), generator=g)).tolist()
    _fp8_case(cuda_device, lens, 30720, 8, 1, 128, 16)
