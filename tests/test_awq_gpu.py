"""GPU parity for the AWQ-int4 path: standalone dequantisation (bit-exact) and the fused dequant + tcgen05 GEMM.
Shape grid after sgl-kernel/tests/test_awq_dequant.py:64-109 (rows x packed cols, fp16 / bf16, group = rows)
plus group-128 cases and the cfg-4 (Qwen2.5-32B) TP shard shapes."""
import pytest
import torch

from oracle import quant_oracle as qo
from sglang_npu_b200 import kernels, synth
from tests.util import assert_close_path, cosine

pytestmark = pytest.mark.gpu


def _full_range_words(shape, gen):
    """All 32 bits random (the reference test's randint(0, 2^31-1) never sets bit 31 -> nibble 7 < 8)."""
    return torch.randint(-2**31, 2**31 - 1, shape, generator=gen, dtype=torch.int64).to(torch.int32)


@pytest.mark.parametrize("rows", [128, 256, 1536, 3584])
@pytest.mark.parametrize("cols", [16, 72, 448, 576])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_awq_dequantize_bit_exact_reference_grid(cuda_device, rows, cols, dtype):
    """group_size = rows (one group), as the reference test does."""
    g = torch.Generator().manual_seed(rows + cols)
    qweight = _full_range_words((rows, cols), g)
    qzeros = _full_range_words((1, cols), g)
    scales = torch.rand((1, cols * 8), generator=g).to(dtype)
    out = kernels.awq_dequantize(qweight.to(cuda_device), scales.to(cuda_device), qzeros.to(cuda_device))
    ref = qo.awq_dequantize(qweight, scales, qzeros, rows)
    assert out.dtype == dtype and out.shape == (rows, cols * 8)
    assert torch.equal(out.cpu().view(torch.int16), ref.to(dtype).view(torch.int16))


@pytest.mark.parametrize("K,N,G", [(256, 128, 128), (1024, 1024, 128), (512, 4736 * 8 // 8, 64), (5120, 7168, 128)])
def test_awq_dequantize_grouped(cuda_device, K, N, G):
    qweight, scales, qzeros = synth.awq_weight(K, N, G, torch.float16, seed=K + N)
    out = kernels.awq_dequantize(qweight.to(cuda_device), scales.to(cuda_device), qzeros.to(cuda_device))
    ref = qo.awq_dequantize(qweight, scales, qzeros, G)
    assert torch.equal(out.cpu().view(torch.int16), ref.view(torch.int16))


def _gemm_case(dev, M, K, N, G, dtype, with_bias=False, seed=0):
    g = torch.Generator().manual_seed(seed + M + K + N)
    qweight = _full_range_words((K, N // 8), g)
    qzeros = _full_range_words((K // G, N // 8), g)
    # (q - z) has std 6.52 -> W std ~ 1/sqrt(K): outputs O(1) for unit-variance activations
    scales = ((torch.rand((K // G, N), generator=g) + 0.5) / (6.52 * K ** 0.5)).to(dtype)
    x = torch.randn((M, K), generator=g).to(dtype)
    bias = torch.randn((N,), generator=g).to(dtype) if with_bias else None
    out = kernels.awq_gemm(x.to(dev), qweight.to(dev), scales.to(dev), qzeros.to(dev), bias.to(dev) if with_bias else None)
    torch.cuda.synchronize()
    ref = qo.awq_linear(x, qweight, scales, qzeros, G, bias)
    return out, ref


@pytest.mark.parametrize("M", [1, 5, 16, 17, 64, 100, 128])
@pytest.mark.parametrize("K,N,G", [(128, 128, 128), (512, 256, 128), (4096, 1024, 128), (2560, 5120, 128), (1024, 160, 64)])
def test_awq_gemm_decode_sizes(cuda_device, M, K, N, G):
    """Weight-streaming (cluster split-K) regime; north_star tolerance on O(1) outputs."""
    out, ref = _gemm_case(cuda_device, M, K, N, G, torch.float16, with_bias=(M == 5))
    assert_close_path(out, ref, f"awq_gemm M={M} K={K} N={N} g={G}")


@pytest.mark.parametrize("M,K,N", [(129, 512, 256), (300, 1024, 1152), (1024, 5120, 1024), (2048, 1024, 3584)])
def test_awq_gemm_prefill_sizes(cuda_device, M, K, N):
    out, ref = _gemm_case(cuda_device, M, K, N, 128, torch.float16, with_bias=(M == 300))
    assert_close_path(out, ref, f"awq_gemm M={M} K={K} N={N}")


@pytest.mark.parametrize("M", [8, 200])
def test_awq_gemm_bf16(cuda_device, M):
    out, ref = _gemm_case(cuda_device, M, 1024, 512, 128, torch.bfloat16)
    # bf16 weights/outputs carry 8 mantissa bits: tolerance scaled to the output peak (2^-7 relative)
    peak = ref.float().abs().max().item()
    assert_close_path(out, ref, f"awq_gemm bf16 M={M}", max_abs=max(2e-2, peak * 2.0 ** -6))


def test_awq_gemm_group_equals_k_and_small_group(cuda_device):
    for (K, G) in [(512, 512), (256, 32)]:
        out, ref = _gemm_case(cuda_device, 33, K, 256, G, torch.float16, seed=G)
        assert_close_path(out, ref, f"awq_gemm group {G}")


@pytest.mark.parametrize("tp", [2, 8])
def test_awq_gemm_cfg4_shard_shapes(cuda_device, tp):
    """Qwen2.5-32B (cfg 4) per-GPU shapes at decode bs=128: qkv (column) and down_proj (row) shards."""
    hidden, inter, hq, hkv, d = 5120, 27648, 40, 8, 128
    n_qkv = (hq // tp + 2 * max(1, hkv // tp)) * d
    for (K, N) in [(hidden, n_qkv), (inter // tp, hidden)]:
        out, ref = _gemm_case(cuda_device, 128, K, N, 128, torch.float16, seed=tp)
        assert_close_path(out, ref, f"awq_gemm cfg4 tp={tp} K={K} N={N}")


def test_awq_linear_method_matches_reference_flow(cuda_device):
    """B200AWQLinearMethod.apply == awq_dequantize -> matmul -> bias (awq.py:401-418)."""
    from sglang_npu_b200.srt.layers.linear import ColumnParallelLinear
    from sglang_npu_b200.srt.layers.quantization import get_quantization_config

    cfg = get_quantization_config("b200_awq").from_config({"w_bit": 4, "q_group_size": 128, "zero_point": True})
    K, N, M = 1024, 768, 48
    with torch.device(cuda_device):
        lin = ColumnParallelLinear(K, N, bias=True, quant_config=cfg, params_dtype=torch.float16, prefix="l.qkv")
    qweight, scales, qzeros = synth.awq_weight(K, N, 128, torch.float16, seed=5)
    lin.qweight.data.copy_(qweight)
    lin.qzeros.data.copy_(qzeros)
    lin.scales.data.copy_(scales)
    lin.bias.data.copy_(torch.randn(N) * 0.1)
    lin.quant_method.process_weights_after_loading(lin)
    x = synth.activations(M, K, torch.float16, seed=9, scale=1.0)
    out, _ = lin(x.to(cuda_device))
    w = kernels.awq_dequantize(lin.qweight, lin.scales, lin.qzeros)
    flow = torch.matmul(x.to(cuda_device), w) + lin.bias
    ref = qo.awq_linear(x, qweight, scales, qzeros, 128, lin.bias.cpu())
    assert_close_path(out, ref, "awq linear vs oracle")
    assert cosine(out.float().cpu(), flow.float().cpu()) > 0.9999


def test_awq_shape_errors(cuda_device):
    dev = cuda_device
    x = torch.zeros((4, 100), dtype=torch.float16, device=dev)
    with pytest.raises(RuntimeError):  # K not a multiple of the group size
        kernels.awq_gemm(x, torch.zeros((100, 16), dtype=torch.int32, device=dev),
                         torch.zeros((1, 128), dtype=torch.float16, device=dev),
                         torch.zeros((1, 16), dtype=torch.int32, device=dev))


def test_awq_repack_is_a_pure_layout_change(cuda_device):
    """qweight_t holds, per (256-channel tile, 128-k block), a [256 channels x 16 words] image whose words carry 8
    consecutive k of one channel in the AWQ nibble order (16-byte chunks swizzled by (row >> 1) & 3); sz pairs the group's
    scale with (magic + zero point).  Unpacking it in torch gives back every nibble / scale / zero of the checkpoint."""
    K, N, G = 512, 264, 128
    g = torch.Generator().manual_seed(4)
    qweight = _full_range_words((K, N // 8), g)
    qzeros = _full_range_words((K // G, N // 8), g)
    order = torch.tensor([0, 4, 1, 5, 2, 6, 3, 7])
    for dtype, magic in ((torch.float16, 0x6400), (torch.bfloat16, 0x4300)):
        scales = torch.rand((K // G, N), generator=g).to(dtype)
        qwt, sz = kernels.awq_repack(qweight.to(cuda_device), scales.to(cuda_device), qzeros.to(cuda_device))
        qwt, sz = qwt.cpu(), sz.cpu()
        nib = lambda words, pos: (words.unsqueeze(-1) >> (4 * pos)) & 0xF  # noqa: E731
        q_ref = nib(qweight, order).reshape(K, N)           # logical [K, N] nibbles of the checkpoint layout
        tiles, nkb = (N + 255) // 256, K // 128
        img = qwt.view(tiles, nkb, 256, 4, 4)               # [tile, kb, row, swizzled chunk, word]
        rows = torch.arange(256)
        src_chunk = torch.arange(4).view(1, 4) ^ ((rows >> 1) & 3).view(256, 1)    # logical chunk c sits at c ^ swizzle
        img = torch.gather(img, 3, src_chunk.view(1, 1, 256, 4, 1).expand(tiles, nkb, 256, 4, 4))
        words = img.reshape(tiles, nkb, 256, 16).permute(0, 2, 1, 3).reshape(tiles * 256, K // 8)  # [channel, word j]
        assert not words[N:].any(), "padding rows past N must be zero"
        q_new = nib(words[:N], order).reshape(N, K).t()      # word (n, j), nibble ORDER[e] = k 8j + e
        assert torch.equal(q_ref, q_new)
        z_ref = nib(qzeros, order).reshape(K // G, N)
        assert torch.equal(sz & 0xFFFF, scales.view(torch.int16).to(torch.int32) & 0xFFFF)
        assert torch.equal((sz >> 16) & 0xFFFF, z_ref.to(torch.int32) | magic)


@pytest.mark.parametrize("M,K,N", [(128, 8192, 256), (64, 16384, 512), (16, 4096, 264)])
def test_awq_gemm_streamk_many_contributors(cuda_device, M, K, N):
    """Few channel tiles, long K: every CTA holds a slice of ONE tile, so the owner sums dozens of partial accumulators
    (the stream-K fix-up path); repeated to show the flag workspace is left clean."""
    for rep in range(3):
        out, ref = _gemm_case(cuda_device, M, K, N, 128, torch.float16, seed=rep)
        assert_close_path(out, ref, f"awq_gemm stream-K M={M} K={K} N={N} rep {rep}")


def test_awq_gemm_is_deterministic(cuda_device):
    g = torch.Generator().manual_seed(1)
    K, N, M = 5120, 3584, 128
    qweight, qzeros = _full_range_words((K, N // 8), g).to(cuda_device), _full_range_words((K // 128, N // 8), g).to(cuda_device)
    scales = (torch.rand((K // 128, N), generator=g) / 400).half().to(cuda_device)
    x = torch.randn((M, K), generator=g).half().to(cuda_device)
    qwt, sz = kernels.awq_repack(qweight, scales, qzeros)
    a = kernels.awq_gemm_repacked(x, qwt, sz, 128).clone()
    for _ in range(3):
        assert torch.equal(kernels.awq_gemm_repacked(x, qwt, sz, 128), a)


def test_awq_gemm_tile_aligned_grid_cfg4_tp4(cuda_device):
    """Above 64 tokens and with no more 256-channel tiles than SMs the grid is tiles x P CTAs (awq.cu run_awq_gemm): every
    tile is cut into exactly P contiguous K ranges.  The four per-GPU GEMMs of cfg 4 at TP 4 give P = 10..14 (7 tiles),
    5 (20 tiles, 10 K blocks), 2 (54 tiles) and 7 with an uneven 54 / 7 split; twice each (flags reused)."""
    for (K, N) in [(5120, 1792), (1280, 5120), (5120, 13824), (6912, 5120)]:
        for rep in range(2):
            out, ref = _gemm_case(cuda_device, 128, K, N, 128, torch.float16, seed=rep)
            assert_close_path(out, ref, f"awq_gemm aligned grid K={K} N={N} rep {rep}")


@pytest.mark.parametrize("per_tile", [0, 1, 3, 7])
def test_awq_gemm_forced_ctas_per_tile_same_bits(cuda_device, per_tile):
    """B200_AWQ_CTAS_PER_TILE (read once per process, hence the subprocess): whatever the K split, partials are summed in
    CTA order, so the result must agree with the oracle — and P = 1 (no fix-up at all) is the yardstick the others are
    held to within fp32 re-association of the K sum."""
    import os
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import sys, torch; sys.path.insert(0, %r)\n"
        "from tests.test_awq_gpu import _gemm_case\n"
        "from tests.util import assert_close_path\n"
        "for (K, N) in [(3456, 5120), (5120, 896)]:\n"
        "    out, ref = _gemm_case(torch.device('cuda:0'), 128, K, N, 128, torch.float16, seed=5)\n"
        "    assert_close_path(out, ref, f'awq forced split K={K} N={N}')\n"
        "print('ok')\n" % (root,))
    env = dict(os.environ, B200_AWQ_CTAS_PER_TILE=str(per_tile))
    res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, cwd=root, timeout=600)
    assert res.returncode == 0 and "ok" in res.stdout, res.stdout[-2000:] + res.stderr[-2000:]
