"""oracle/quant_oracle.py held to fixtures produced by the REFERENCE'S OWN known-answer functions
(tests/golden/make_golden_quant.py lifts them out of /root/reference with `ast` and runs them on seeded inputs).
This is what pins the FP8-block / FP8-channel / AWQ / int8 oracle."""
import os

import pytest
import torch

from oracle import quant_oracle as qo

GOLD = os.path.join(os.path.dirname(__file__), "golden")
DT = {"torch.bfloat16": torch.bfloat16, "torch.float16": torch.float16, "torch.float32": torch.float32}


def load(name):
    return torch.load(os.path.join(GOLD, name + ".pt"), weights_only=True)


def f8(u8):
    return u8.view(torch.float8_e4m3fn)


def test_group_quant_fp8_matches_reference_function():
    """Bit-exact codes and scales, with ONE stated exception: an all-zero group.  The reference's TEST function
    clamps eps in the input dtype (bf16 rounds 1e-10 to 1.0012e-10, fp16 flushes it to 0 -> NaN codes), its KERNELS
    clamp in fp32 (fp8_kernel.py:145-150, per_token_group_quant_8bit.cu); the oracle follows the kernels."""
    for c in load("quant_fp8_group"):
        x, g = c["x"], c["group"]
        q, s = qo.per_token_group_quant_fp8(x.contiguous(), g)
        zero_group = (x.reshape(-1, g).abs().amax(dim=-1) == 0).reshape(s.shape)
        assert zero_group.any(), "fixture must contain the all-zero-group case"
        ok = ~zero_group
        assert torch.equal(s[ok], c["s"][ok]), c["dtype"]
        okq = ok.repeat_interleave(g, dim=-1)
        assert torch.equal(q.view(torch.uint8)[okq], c["q"][okq]), c["dtype"]
        # the exception, explicitly: oracle scale = 1e-10/448 (fp32 clamp) and zero codes
        assert torch.all(s[zero_group] == torch.tensor(1e-10, dtype=torch.float32) / 448.0)
        assert torch.all(q.view(torch.uint8)[zero_group.repeat_interleave(g, dim=-1)] == 0)
        ref_s = c["s"][zero_group]
        if c["dtype"] == "torch.float32":
            assert torch.equal(ref_s, s[zero_group])  # fp32 inputs: no deviation at all
        elif c["dtype"] == "torch.float16":
            assert torch.all(ref_s == 0)  # the test function's fp16 clamp flushes eps to zero
        else:
            assert torch.all((ref_s - s[zero_group]).abs() / s[zero_group] < 2e-3)  # bf16 rounding of eps


def test_block_fp8_matmul_matches_both_reference_functions():
    for c in load("quant_fp8_block_mm"):
        dt = DT[c["out_dtype"]]
        out = qo.w8a8_block_fp8_matmul(f8(c["a"]), f8(c["b"]), c["sa"], c["sb"], (128, 128), dt)
        ref = c["out_native"]
        # same per-K-block fp32 partial products; only the fp32 matmul summation order may differ (threads)
        tol = dict(rtol=2e-5, atol=2e-5 * ref.float().abs().max().item()) if dt == torch.float32 else dict(rtol=0, atol=0)
        if dt == torch.float32:
            torch.testing.assert_close(out, ref, **tol)
        else:  # at most one output-dtype ulp on a handful of elements
            d = (out.float() - ref.float()).abs()
            ulp = torch.finfo(dt).eps * ref.float().abs().clamp_min(1e-30)
            assert (d <= ulp).all() and (d > 0).float().mean() < 0.02
        # baseline_scaled_mm scales the operands BEFORE one fp32 matmul: equal up to fp32 rounding
        d = (out.float() - c["out_baseline"].float()).abs().max().item()
        assert d <= 2e-2 * max(1.0, ref.float().abs().max().item())


def test_channel_fp8_matches_reference_functions():
    for c in load("quant_fp8_chan"):
        if c["kind"] == "mm":
            dt = DT[c["out_dtype"]]
            out = qo.fp8_scaled_mm(f8(c["a"]), f8(c["b"]), c["sa"], c["sb"], dt, c["bias"])
            ref = c["out"]
            # the reference test rounds to out_dtype BEFORE adding the bias (test_fp8_gemm.py:11-13); the kernel and
            # the oracle add in fp32 — one extra output rounding is the whole difference
            d = (out.float() - ref.float()).abs()
            mag = ref.float().abs()
            if c["bias"] is not None:  # cancellation: the rounding that differs happened at the pre-bias magnitude
                mag = torch.maximum(mag, c["bias"].float().abs().view(1, -1))
            assert (d <= 2 * torch.finfo(dt).eps * mag + 1e-30).all()
        else:
            q, s = qo.per_token_quant_fp8(c["x"])
            assert torch.equal(s.view(-1), c["s"])
            assert torch.equal(q.view(torch.uint8), c["q"])


def test_awq_dequant_matches_reference_function():
    for c in load("quant_awq"):
        w = qo.awq_dequantize(c["qweight"], c["scales"], c["qzeros"], c["group"])
        assert w.dtype == c["w"].dtype
        assert torch.equal(w.view(torch.int16), c["w"].view(torch.int16))


def test_int8_matches_reference_functions():
    for c in load("quant_int8"):
        xq, xs = qo.per_token_quant_int8(c["x"])
        assert torch.equal(xq, c["xq"]) and torch.equal(xs, c["xs"])
        out = qo.int8_scaled_mm(xq, c["w"], xs, c["ws"], torch.bfloat16, c["bias"])
        for key in ("out_native", "out_scaled_mm"):
            d = (out.float() - c[key].float()).abs()
            ulp = torch.finfo(torch.bfloat16).eps * c[key].float().abs().clamp_min(1e-30)
            assert (d <= ulp).all(), key
        lin = qo.w8a8_int8_linear(c["x"], c["w"], c["ws"].view(-1, 1), c["bias"])
        assert torch.equal(lin, out)
