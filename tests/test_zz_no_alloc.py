"""b200_argmax_static: the greedy argmax with its scratch in a file-scope `static __device__` array of the library
(no cudaMalloc / cudaFree inside the call; 4096 rows at most), checked through the C-ABI against the existing entry
point b200_argmax (caller-owned workspace) — index work, so the bar is equality.  Collected last on purpose;
passes alone in a fresh process and after the other tests (it shares nothing with them:
the scratch is cleared on the stream before every launch)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_ROWS = 4096  # kArgmaxStaticMaxRows in csrc/elementwise.cu


def _both(lib, logits, stream):
    from sglang_npu_b200 import kernels
    from sglang_npu_b200._cabi import dtype_code

    rows, V = logits.shape
    dev = logits.device
    ws = torch.zeros(max(rows, 1) * 16, dtype=torch.uint8, device=dev)
    want = torch.full((rows,), -1, dtype=torch.int64, device=dev)
    got = torch.full((rows,), -2, dtype=torch.int64, device=dev)
    torch.cuda.synchronize(dev)
    kernels.check(lib.b200_argmax(want.data_ptr(), logits.data_ptr(), rows, V, logits.stride(0), dtype_code(logits.dtype),
                                  ws.data_ptr(), ws.numel(), stream), "b200_argmax")
    kernels.check(lib.b200_argmax_static(got.data_ptr(), logits.data_ptr(), rows, V, logits.stride(0),
                                         dtype_code(logits.dtype), stream), "b200_argmax_static")
    torch.cuda.synchronize(dev)
    return got, want


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_argmax_static_matches_argmax_on_two_calls_in_a_row(cuda_device, dtype):
    from sglang_npu_b200 import _cabi

    lib = _cabi.load()
    dev = cuda_device
    stream = torch.cuda.current_stream(dev).cuda_stream
    g = torch.Generator(device=dev).manual_seed(11)
    # two calls in a row over the SAME scratch, different shapes: a decode batch over the Llama vocabulary (every row
    # split over several CTAs -> keys + counters in the scratch), then 100 rows (3 CTAs per row) over a ragged vocabulary with ties
    first = torch.randn((64, 128256), generator=g, device=dev).to(dtype)
    got, want = _both(lib, first, stream)
    assert torch.equal(got, want)
    assert torch.equal(got, first.float().argmax(-1))
    second = torch.randn((100, 32003 + 5), generator=g, device=dev).to(dtype)[:, :32003]  # row stride 32008, V odd
    second[:, 17] = 9.0
    second[:, 31999] = 9.0  # equal maxima: the lowest index wins
    got2, want2 = _both(lib, second, stream)
    assert torch.equal(got2, want2)
    assert bool((got2 == 17).all())
    # and the first input again: nothing of the second call is left behind
    got3, _ = _both(lib, first, stream)
    assert torch.equal(got3, want)


def test_argmax_static_limits(cuda_device):
    from sglang_npu_b200 import _cabi
    from sglang_npu_b200._cabi import dtype_code

    lib = _cabi.load()
    dev = cuda_device
    stream = torch.cuda.current_stream(dev).cuda_stream
    x = torch.randn((MAX_ROWS, 4096), device=dev).to(torch.bfloat16)
    got, want = _both(lib, x, stream)  # exactly the maximum is accepted
    assert torch.equal(got, want)
    big = torch.zeros((MAX_ROWS + 1, 64), dtype=torch.bfloat16, device=dev)
    out = torch.zeros(MAX_ROWS + 1, dtype=torch.int64, device=dev)
    rc = lib.b200_argmax_static(out.data_ptr(), big.data_ptr(), MAX_ROWS + 1, 64, 64, dtype_code(torch.bfloat16), stream)
    assert rc != 0
    msg = lib.b200_last_error().decode()
    assert "4096" in msg and "argmax_static" in msg, msg
    assert lib.b200_argmax_static(out.data_ptr(), big.data_ptr(), 0, 64, 64, dtype_code(torch.bfloat16), stream) == 0
    torch.cuda.synchronize(dev)
