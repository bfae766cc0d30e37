"""Seeded synthetic weights for the FP8 / AWQ linear methods (used by layer_stack.load_synthetic_weights);
distributions follow the reference tests (sgl-kernel/tests/test_fp8_blockwise_gemm.py:66-75,
test_awq_dequant.py:81-101) rescaled so that activations stay O(1) through a deep stack."""
from __future__ import annotations

import math

import torch


def _load(lin, param, tensor, sid):
    if sid is not None:
        lin.weight_loader(param, tensor, sid)
    else:
        lin.weight_loader(param, tensor)


def fill(lin, sid, n: int, k: int, gen: torch.Generator, device, gain: float) -> None:
    method = type(lin.quant_method).__name__
    if method == "B200Fp8LinearMethod":
        w = ((torch.rand((n, k), generator=gen, device=device) - 0.5) * 2 * 448.0).to(torch.float8_e4m3fn)
        # effective std of U(-448,448) is 258.6; block scales ~ |N(0,1)| like the reference test, mean-normalised
        s = (torch.randn((n // 128, k // 128), generator=gen, device=device).abs() + 0.5) * (
            gain / (258.6 * 1.3 * math.sqrt(k)))
        _load(lin, lin.weight, w, sid)
        _load(lin, lin.weight_scale_inv, s, sid)
    elif method == "B200W8A8Fp8LinearMethod":
        w = ((torch.rand((n, k), generator=gen, device=device) - 0.5) * 2 * 448.0).to(torch.float8_e4m3fn)
        s = (torch.rand((n, 1), generator=gen, device=device) + 0.5) * (gain / (258.6 * math.sqrt(k)))
        _load(lin, lin.weight, w, sid)
        _load(lin, lin.weight_scale, s, sid)
    elif method == "B200AWQLinearMethod":
        g = lin.quant_method.quant_config.group_size
        qweight = torch.randint(0, 2**31 - 1, (k, n // 8), generator=gen, device=device, dtype=torch.int64).to(torch.int32)
        qzeros = torch.randint(0, 2**31 - 1, (k // g, n // 8), generator=gen, device=device, dtype=torch.int64).to(torch.int32)
        # (q - z) of two U{0..15} has std 6.52
        scales = ((torch.rand((k // g, n), generator=gen, device=device) + 0.5) * (gain / (6.52 * math.sqrt(k)))).to(
            lin.params_dtype)
        _load(lin, lin.qweight, qweight, sid)
        _load(lin, lin.qzeros, qzeros, sid)
        _load(lin, lin.scales, scales, sid)
    else:
        raise NotImplementedError(method)
