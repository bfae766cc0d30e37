"""GPU parity for the W8A8-int8 linear path: per-token quant codes and int32 accumulators bit-exact,
bf16 outputs within the north_star tolerance (shape grid after sgl-kernel/tests/test_int8_gemm.py:36-42
and test/srt/cpu/test_gemm.py:83-126)."""
import pytest
import torch

from oracle import quant_oracle as qo
from sglang_npu_b200 import kernels, synth
from tests.util import assert_close_path

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("M,K", [(1, 512), (7, 896), (64, 4096), (33, 14336), (300, 1024)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_per_token_quant_int8_bit_exact(cuda_device, M, K, dtype):
    x = synth.activations(M, K, dtype, seed=M + K)
    x[0, :8] = 0  # exercise exact zeros / ties
    if M > 2:
        x[2] = 0  # all-zero row -> absmax clamp 1e-10
    xq, xs = kernels.per_token_quant_int8(x.to(cuda_device))
    rq, rs = qo.per_token_quant_int8(x)
    assert torch.equal(xq.cpu(), rq)
    assert torch.equal(xs.cpu().view(torch.int32), rs.view(torch.int32))


def _run_mm(dev, M, N, K, with_bias, out_dtype, seed=0, cluster_splitk=False):
    g = torch.Generator().manual_seed(seed + M * 7 + N * 3 + K)
    a = torch.randint(-127, 128, (M, K), generator=g, dtype=torch.int32).to(torch.int8)
    w = torch.randint(-127, 128, (N, K), generator=g, dtype=torch.int32).to(torch.int8)
    sa = torch.rand((M,), generator=g) * 1e-2
    sb = torch.rand((N,), generator=g) * 1e-2
    bias = (torch.randn((N,), generator=g) * 10).to(out_dtype) if with_bias else None
    out, acc = kernels.int8_scaled_mm(a.to(dev), w.to(dev).t(), sa.to(dev), sb.to(dev), out_dtype,
                                      bias.to(dev) if with_bias else None, return_acc=True,
                                      cluster_splitk=cluster_splitk)
    torch.cuda.synchronize()
    ref_acc = qo.int8_accumulate(a, w)
    ref = qo.int8_scaled_mm(a, w, sa, sb, out_dtype, bias)
    return out.cpu(), acc.cpu(), ref, ref_acc


# decode (weight-streaming, split-K) shapes incl. ragged M / N / K tails
@pytest.mark.parametrize("M", [1, 5, 16, 17, 64, 100, 128, 256])
@pytest.mark.parametrize("N,K", [(16, 512), (1152, 896), (896, 4864), (6144, 4096), (4096, 14336), (136, 144)])
def test_int8_scaled_mm_decode_shapes(cuda_device, M, N, K):
    out, acc, ref, ref_acc = _run_mm(cuda_device, M, N, K, with_bias=(M % 2 == 1), out_dtype=torch.bfloat16)
    assert torch.equal(acc, ref_acc), "int32 accumulators must be bit-exact"
    # fp32 epilogue in the oracle's order, one rounding to bf16: bit-exact as well
    assert torch.equal(out.view(torch.int16), ref.view(torch.int16))


# the cluster split-K kernel (no workspace) must give the same bits as the persistent stream-K one
@pytest.mark.parametrize("M", [1, 17, 64, 256])
@pytest.mark.parametrize("N,K", [(1152, 896), (6144, 4096), (4096, 14336), (136, 144)])
def test_int8_scaled_mm_cluster_splitk(cuda_device, M, N, K):
    out, acc, ref, ref_acc = _run_mm(cuda_device, M, N, K, with_bias=(M % 2 == 1), out_dtype=torch.bfloat16,
                                     cluster_splitk=True)
    assert torch.equal(acc, ref_acc)
    assert torch.equal(out.view(torch.int16), ref.view(torch.int16))


def test_int8_stream_k_small_token_tiles_subprocess(cuda_device):
    """The stream-K kernel is dispatched from 33 tokens up; its 16 / 32-token tiles are reachable with
    B200_GEMM_STREAMK_MIN_M=1 (read once per process), so they are checked in a child process."""
    import os
    import subprocess
    import sys
    code = (
        "import torch\n"
        "from oracle import quant_oracle as qo\n"
        "from sglang_npu_b200 import kernels\n"
        "g = torch.Generator().manual_seed(11)\n"
        "for (M, N, K) in [(1, 1152, 896), (7, 6144, 4096), (16, 4096, 14336), (32, 136, 144), (19, 28672, 4096)]:\n"
        "    a = torch.randint(-127, 128, (M, K), generator=g, dtype=torch.int32).to(torch.int8)\n"
        "    w = torch.randint(-127, 128, (N, K), generator=g, dtype=torch.int32).to(torch.int8)\n"
        "    sa, sb = torch.rand(M, generator=g), torch.rand(N, generator=g)\n"
        "    out, acc = kernels.int8_scaled_mm(a.cuda(), w.cuda().t(), sa.cuda(), sb.cuda(), torch.bfloat16, return_acc=True)\n"
        "    assert torch.equal(acc.cpu(), qo.int8_accumulate(a, w)), (M, N, K)\n"
        "    ref = qo.int8_scaled_mm(a, w, sa, sb, torch.bfloat16, None)\n"
        "    assert torch.equal(out.cpu().view(torch.int16), ref.view(torch.int16)), (M, N, K)\n"
        "print('ok')\n"
    )
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, B200_GEMM_STREAMK_MIN_M="1", PYTHONPATH=root)
    res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, cwd=root, timeout=600)
    assert res.returncode == 0 and "ok" in res.stdout, res.stdout + res.stderr


def test_int8_stream_k_many_launches_no_drift(cuda_device):
    """Stream-K flags are reset by their consumer: 200 back-to-back launches over shapes with different CTA
    ranges must keep producing the same bits (a stale flag would let an owner read an old partial)."""
    dev = cuda_device
    g = torch.Generator().manual_seed(5)
    shapes = [(64, 6144, 4096), (64, 4096, 14336), (33, 1152, 896), (64, 28672, 4096)]
    ops = []
    for (M, N, K) in shapes:
        a = torch.randint(-127, 128, (M, K), generator=g, dtype=torch.int32).to(torch.int8)
        w = torch.randint(-127, 128, (N, K), generator=g, dtype=torch.int32).to(torch.int8)
        ops.append((a.to(dev), w.to(dev), torch.ones(M, device=dev), torch.ones(N, device=dev), qo.int8_accumulate(a, w)))
    accs = []
    for i in range(200):
        a, w, sa, sb, _ = ops[i % len(ops)]
        accs.append(kernels.int8_scaled_mm(a, w.t(), sa, sb, torch.bfloat16, return_acc=True)[1])
    torch.cuda.synchronize()
    for i, acc in enumerate(accs):
        assert torch.equal(acc.cpu(), ops[i % len(ops)][4]), f"launch {i}"


@pytest.mark.parametrize("M", [257, 512, 1000, 2048])
@pytest.mark.parametrize("N,K", [(128, 512), (6144, 4096), (264, 1040), (28672, 4096)])
def test_int8_scaled_mm_prefill_shapes(cuda_device, M, N, K):
    out, acc, ref, ref_acc = _run_mm(cuda_device, M, N, K, with_bias=(N == 264), out_dtype=torch.float16)
    assert torch.equal(acc, ref_acc), "int32 accumulators must be bit-exact"
    assert torch.equal(out.view(torch.int16), ref.view(torch.int16))


# TP shard shapes (Llama-3.1-8B at tp 4 / tp 8) at the decode batches of the weak-scaling bench (64 tokens per GPU): the
# stream-K kernel cuts the batch into 64-token blocks, incl. a ragged last block and 256 < M <= 512
@pytest.mark.parametrize("M", [65, 200, 256, 300, 512])
@pytest.mark.parametrize("N,K", [(1536, 4096), (4096, 1024), (7168, 4096), (4096, 3584), (768, 4096), (4096, 512)])
def test_int8_scaled_mm_token_blocks_tp_shards(cuda_device, M, N, K):
    out, acc, ref, ref_acc = _run_mm(cuda_device, M, N, K, with_bias=(M == 200), out_dtype=torch.bfloat16)
    assert torch.equal(acc, ref_acc), "int32 accumulators must be bit-exact"
    assert torch.equal(out.view(torch.int16), ref.view(torch.int16))


@pytest.mark.parametrize("bn", [128, 256])
def test_int8_stream_k_wider_token_blocks_subprocess(cuda_device, bn):
    """B200_SK_BN (read once per process) selects 128- / 256-token blocks instead of 64: same bits."""
    import os
    import subprocess
    import sys
    code = (
        "import torch\n"
        "from oracle import quant_oracle as qo\n"
        "from sglang_npu_b200 import kernels\n"
        "g = torch.Generator().manual_seed(3)\n"
        "for (M, N, K) in [(130, 1536, 4096), (256, 4096, 1024), (300, 7168, 4096), (512, 768, 4096), (500, 4096, 3584)]:\n"
        "    a = torch.randint(-127, 128, (M, K), generator=g, dtype=torch.int32).to(torch.int8)\n"
        "    w = torch.randint(-127, 128, (N, K), generator=g, dtype=torch.int32).to(torch.int8)\n"
        "    sa, sb = torch.rand(M, generator=g), torch.rand(N, generator=g)\n"
        "    out, acc = kernels.int8_scaled_mm(a.cuda(), w.cuda().t(), sa.cuda(), sb.cuda(), torch.bfloat16, return_acc=True)\n"
        "    assert torch.equal(acc.cpu(), qo.int8_accumulate(a, w)), (M, N, K)\n"
        "    ref = qo.int8_scaled_mm(a, w, sa, sb, torch.bfloat16, None)\n"
        "    assert torch.equal(out.cpu().view(torch.int16), ref.view(torch.int16)), (M, N, K)\n"
        "print('ok')\n"
    )
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, B200_SK_BN=str(bn), PYTHONPATH=root)
    res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, cwd=root, timeout=600)
    assert res.returncode == 0 and "ok" in res.stdout, res.stdout + res.stderr


def test_int8_split_k_workspace_is_self_cleaning(cuda_device):
    """Back-to-back calls reuse the split-K workspace; results must not drift."""
    for i in range(3):
        out, acc, ref, ref_acc = _run_mm(cuda_device, 64, 4096, 4096, False, torch.bfloat16, seed=i)
        assert torch.equal(acc, ref_acc)


def test_int8_argument_errors(cuda_device):
    dev = cuda_device
    a = torch.zeros((4, 24), dtype=torch.int8, device=dev)
    w = torch.zeros((16, 24), dtype=torch.int8, device=dev)
    s = torch.ones(4, device=dev)
    with pytest.raises(RuntimeError, match="multiple of 16"):
        kernels.int8_scaled_mm(a, w.t(), s, torch.ones(16, device=dev), torch.bfloat16)
    a = torch.zeros((4, 32), dtype=torch.int8, device=dev)
    with pytest.raises(RuntimeError, match="column-major"):
        kernels.int8_scaled_mm(a, torch.zeros((32, 16), dtype=torch.int8, device=dev), s, torch.ones(16, device=dev),
                               torch.bfloat16)
    with pytest.raises(NotImplementedError):
        kernels.int8_scaled_mm(a, torch.zeros((16, 32), dtype=torch.int8, device=dev).t(), s,
                               torch.ones(16, device=dev), torch.float32)
