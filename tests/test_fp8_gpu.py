"""GPU parity for the FP8 linear paths (block-scaled and per-channel); shape grid after
sgl-kernel/tests/test_fp8_blockwise_gemm.py:76-82 and python/sglang/test/test_block_fp8.py."""
import pytest
import torch

from oracle import quant_oracle as qo
from sglang_npu_b200 import kernels, synth
from tests.util import assert_close_path, cosine

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("M,K", [(1, 128), (7, 896), (64, 4096), (83, 14336)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("col_major", [False, True])
def test_per_token_group_quant_fp8_bit_exact(cuda_device, M, K, dtype, col_major):
    x = synth.activations(M, K, dtype, seed=M + K, scale=1.0)
    x[0, :128] = 0  # an all-zero group -> eps clamp
    xq, xs = kernels.sgl_per_token_group_quant_fp8(x.to(cuda_device), 128, column_major_scales=col_major)
    rq, rs = qo.per_token_group_quant_fp8(x, 128)
    assert torch.equal(xs.cpu().contiguous().view(torch.int32), rs.view(torch.int32))
    assert torch.equal(xq.cpu().view(torch.uint8), rq.view(torch.uint8))
    if col_major:
        assert xs.stride(0) == 1  # "TMA-aligned" column-major scale layout (fp8_utils.py:203-205)


def test_per_token_quant_fp8(cuda_device):
    x = synth.activations(33, 4096, torch.bfloat16, seed=3, scale=2.0)
    xq, xs = kernels.sgl_per_token_quant_fp8(x.to(cuda_device))
    rq, rs = qo.per_token_quant_fp8(x)
    assert torch.equal(xs.cpu().view(torch.int32), rs.view(torch.int32))
    # reference test tolerance (sgl-kernel/tests/test_per_token_quant_fp8.py:51-53)
    torch.testing.assert_close(xq.cpu().float(), rq.float(), rtol=1e-3, atol=1e-3)


def _block_case(dev, M, N, K, out_dtype, with_bias=False, seed=0):
    g = torch.Generator().manual_seed(seed + M + N + K)
    a = ((torch.rand((M, K), generator=g) - 0.5) * 2 * 448).to(torch.float8_e4m3fn)
    w = ((torch.rand((N, K), generator=g) - 0.5) * 2 * 448).to(torch.float8_e4m3fn)
    sa = torch.randn((M, K // 128), generator=g).abs() * 1e-3 + 1e-4
    sb = torch.randn((N // 128, K // 128), generator=g).abs() * 1e-3 + 1e-4
    bias = torch.randn((N,), generator=g).to(out_dtype) if with_bias else None
    out = kernels.fp8_blockwise_scaled_mm(a.to(dev), w.to(dev).t(), sa.to(dev), sb.to(dev).t(), out_dtype,
                                          bias.to(dev) if with_bias else None)
    torch.cuda.synchronize()
    ref = qo.w8a8_block_fp8_matmul(a, w, sa, sb, (128, 128), torch.float32)
    if with_bias:
        ref = ref + bias.float()
    return out.cpu().float(), ref


@pytest.mark.parametrize("M", [1, 3, 64, 127, 128, 129, 300, 1024])
@pytest.mark.parametrize("N,K", [(128, 128), (512, 1024), (6144, 4096), (4096, 14336)])
def test_fp8_blockwise_scaled_mm(cuda_device, M, N, K):
    out, ref = _block_case(cuda_device, M, N, K, torch.bfloat16, with_bias=(M == 3))
    # reference tolerance: rtol 0.02 / atol 1 on outputs of magnitude ~ 1e2 (test_fp8_blockwise_gemm.py:80-82);
    # tighter here: bf16 rounding of the fp32 result + fp32 summation-order noise
    scale = ref.abs().max().item()
    assert (out - ref).abs().max().item() <= 1e-2 * scale + 1e-3
    assert cosine(out, ref) > 0.9999


@pytest.mark.parametrize("M", [5, 64, 200, 700])
def test_block_fp8_linear_end_to_end(cuda_device, M):
    """Fp8LinearMethod.apply: quant + GEMM on realistic activations, north_star tolerance."""
    N, K = 1024, 2048
    x = synth.activations(M, K, torch.bfloat16, seed=M, scale=1.0)
    w, ws = synth.fp8_block_weight(N, K, seed=2)
    ws = ws * 30  # outputs O(1)
    xq, xs = kernels.sgl_per_token_group_quant_fp8(x.to(cuda_device), 128)
    out = kernels.fp8_blockwise_scaled_mm(xq, w.to(cuda_device).t(), xs, ws.to(cuda_device).t(), torch.bfloat16)
    ref = qo.w8a8_block_fp8_linear(x, w, ws)
    peak = ref.float().abs().max().item()
    assert_close_path(out, ref, f"block fp8 linear M={M}", max_abs=max(2e-2, peak * 2.0 ** -7))


@pytest.mark.parametrize("M,N,K", [(1, 256, 512), (64, 6144, 4096), (130, 1024, 1024), (512, 4096, 4096)])
def test_fp8_scaled_mm_per_channel(cuda_device, M, N, K):
    g = torch.Generator().manual_seed(M + N)
    a = ((torch.rand((M, K), generator=g) - 0.5) * 2 * 448).to(torch.float8_e4m3fn)
    w = ((torch.rand((N, K), generator=g) - 0.5) * 2 * 448).to(torch.float8_e4m3fn)
    sa = torch.rand((M,), generator=g) * 1e-3
    sb = torch.rand((N,), generator=g) * 1e-3
    bias = torch.randn((N,), generator=g).to(torch.bfloat16)
    dev = cuda_device
    out = kernels.fp8_scaled_mm(a.to(dev), w.to(dev).t(), sa.to(dev), sb.to(dev), torch.bfloat16, bias.to(dev))
    ref = qo.fp8_scaled_mm(a, w, sa, sb, torch.float32, bias)
    scale = ref.abs().max().item()
    assert (out.cpu().float() - ref).abs().max().item() <= 1e-2 * scale + 1e-3
    assert cosine(out.cpu().float(), ref) > 0.9999


def test_fp8_shape_errors(cuda_device):
    dev = cuda_device
    a = torch.zeros((4, 192), dtype=torch.float8_e4m3fn, device=dev)
    w = torch.zeros((128, 192), dtype=torch.float8_e4m3fn, device=dev)
    with pytest.raises(RuntimeError):
        kernels.fp8_blockwise_scaled_mm(a, w.t(), torch.ones((4, 1), device=dev), torch.ones((1, 1), device=dev),
                                        torch.bfloat16)
