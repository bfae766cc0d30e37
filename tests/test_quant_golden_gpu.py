"""The CUDA quant kernels against fixtures produced by the REFERENCE'S OWN known-answer functions
(tests/golden/make_golden_quant.py).  No oracle in between: kernel output vs reference-function output."""
import os

import pytest
import torch

from sglang_npu_b200 import kernels

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden")
DT = {"torch.bfloat16": torch.bfloat16, "torch.float16": torch.float16, "torch.float32": torch.float32}


def load(name):
    return torch.load(os.path.join(GOLD, name + ".pt"), weights_only=True)


def f8(u8):
    return u8.view(torch.float8_e4m3fn)


def test_group_quant_fp8_vs_reference_function(cuda_device):
    n = 0
    for c in load("quant_fp8_group"):
        if c["dtype"] == "torch.float32":
            continue  # activations on this path are bf16 / fp16
        x, g = c["x"], c["group"]
        q, s = kernels.sgl_per_token_group_quant_fp8(x.to(cuda_device), g)
        ok = (x.reshape(-1, g).abs().amax(dim=-1) != 0).reshape(c["s"].shape)  # all-zero groups: fp32 eps clamp (kernel rule)
        assert torch.equal(s.cpu()[ok], c["s"][ok])
        okq = ok.repeat_interleave(g, dim=-1)
        assert torch.equal(q.cpu().view(torch.uint8)[okq], c["q"][okq])
        assert torch.all(q.cpu().view(torch.uint8)[~okq] == 0)
        n += 1
    assert n == 4


def test_block_fp8_mm_vs_reference_functions(cuda_device):
    dev = cuda_device
    for c in load("quant_fp8_block_mm"):
        dt = DT[c["out_dtype"]]
        if dt == torch.float32:
            continue
        a, b = f8(c["a"]).to(dev), f8(c["b"]).to(dev)
        out = kernels.fp8_blockwise_scaled_mm(a, b.t(), c["sa"].to(dev), c["sb"].to(dev).t(), dt).cpu().float()
        for key in ("out_native", "out_baseline"):
            ref = c[key].float()
            # the reference's own gate is rtol 0.02 / atol 1 (test_fp8_blockwise_gemm.py:80-82); here: two output ulps of
            # the result's own scale (fp32 accumulation order inside a 128-K block is the only freedom)
            tol = 2 * torch.finfo(dt).eps * ref.abs().clamp_min(ref.abs().max() * 2.0 ** -6)
            assert ((out - ref).abs() <= tol).all(), (key, c["out_dtype"], tuple(out.shape))


def test_channel_fp8_vs_reference_functions(cuda_device):
    dev = cuda_device
    for c in load("quant_fp8_chan"):
        if c["kind"] == "mm":
            dt = DT[c["out_dtype"]]
            bias = None if c["bias"] is None else c["bias"].to(dev)
            out = kernels.fp8_scaled_mm(f8(c["a"]).to(dev), f8(c["b"]).to(dev).t(), c["sa"].to(dev), c["sb"].to(dev), dt,
                                        bias).cpu().float()
            ref = c["out"].float()
            mag = ref.abs()
            if c["bias"] is not None:
                mag = torch.maximum(mag, c["bias"].float().abs().view(1, -1))
            tol = 2 * torch.finfo(dt).eps * mag.clamp_min(ref.abs().max() * 2.0 ** -6)
            assert ((out - ref).abs() <= tol).all()
        else:
            q, s = kernels.sgl_per_token_quant_fp8(c["x"].to(dev))
            assert torch.equal(s.cpu().view(-1), c["s"])
            # reference test tolerance (sgl-kernel/tests/test_per_token_quant_fp8.py:51-53): the kernel multiplies by
            # the reciprocal scale, one e4m3 ulp at most on ties
            torch.testing.assert_close(q.cpu().float(), f8(c["q"]).float(), rtol=0.07, atol=1e-3)
            assert (q.cpu().view(torch.uint8) != c["q"]).float().mean() < 0.01


def test_awq_dequant_vs_reference_function(cuda_device):
    dev = cuda_device
    for c in load("quant_awq"):
        w = kernels.awq_dequantize(c["qweight"].to(dev), c["scales"].to(dev), c["qzeros"].to(dev)).cpu()
        if c["scales"].dtype == torch.float16:
            assert torch.equal(w.view(torch.int16), c["w"].view(torch.int16))  # (q - z) * s in fp16: exact
        else:  # bf16: the reference kernel's own gate (test_awq_dequant.py:107-109)
            torch.testing.assert_close(w.float(), c["w"].float(), rtol=1e-3, atol=1e-5)


def test_awq_gemm_vs_reference_dequant(cuda_device):
    """AWQLinearMethod.apply = awq_dequantize + matmul (awq.py:401-418): fused kernel vs fp32 matmul on the reference
    function's dequantised matrix."""
    dev = cuda_device
    g = torch.Generator().manual_seed(3)
    for c in load("quant_awq"):
        if c["scales"].dtype != torch.float16 or c["group"] != 128:
            continue
        K, N = c["w"].shape
        for M in (1, 33, 128):
            x = (torch.randn(M, K, generator=g) * 0.5).to(torch.float16)
            out = kernels.awq_gemm(x.to(dev), c["qweight"].to(dev), c["scales"].to(dev), c["qzeros"].to(dev)).cpu().float()
            ref = x.float() @ c["w"].float()
            peak = ref.abs().max().item()
            assert (out - ref).abs().max().item() <= peak * 2.0 ** -9 + 1e-3


def test_int8_vs_reference_functions(cuda_device):
    dev = cuda_device
    for c in load("quant_int8"):
        xq, xs = kernels.per_token_quant_int8(c["x"].to(dev))
        assert torch.equal(xq.cpu(), c["xq"]) and torch.equal(xs.cpu(), c["xs"])
        bias = None if c["bias"] is None else c["bias"].to(dev)
        out, acc = kernels.int8_scaled_mm(xq, c["w"].to(dev).t(), xs, c["ws"].to(dev), torch.bfloat16, bias, return_acc=True)
        exact = (c["xq"].double() @ c["w"].double().t()).to(torch.int32)
        assert torch.equal(acc.cpu(), exact)
        for key in ("out_native", "out_scaled_mm"):
            ref = c[key].float()
            assert ((out.cpu().float() - ref).abs() <= torch.finfo(torch.bfloat16).eps * ref.abs() + 1e-5).all(), key
