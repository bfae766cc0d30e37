"""TEST INFRASTRUCTURE (oracle) — CPU restatement of the quantised-linear arithmetic.

Parity: PINNED on the reference's own test restatements (the reference's quant Python cannot be
imported here — SURVEY §8c — and ships no stored vectors): each function below follows the
torch restatement the reference's tests use as THEIR known answer, cited per function.  The int8
path is additionally pinned against the reference's compiled CPU kernel
(oracle/_ref: int8_scaled_mm_with_quant, sgl-kernel/csrc/cpu/gemm_int8.cpp:423-489) in
tests/test_oracle_ref_cpu.py when oracle/_ref is built.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

AWQ_REVERSE_ORDER = [0, 4, 1, 5, 2, 6, 3, 7]


# ---- W8A8 int8 --------------------------------------------------------------------------------
def per_token_quant_int8(x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """test/srt/cpu/utils.py:24-32 (same formula as int8_kernel.py:35-42; torch.round = half-even)."""
    x = x.float()
    absmax = x.abs().max(dim=-1).values
    absmax = absmax.clamp_min(1e-10).unsqueeze(-1)
    scale_x = absmax / 127
    x_q = x.mul(127 / absmax)
    x_q = torch.round(x_q).to(torch.int8)
    return x_q, scale_x


def int8_accumulate(a_q: torch.Tensor, w_q: torch.Tensor) -> torch.Tensor:
    """Exact int32 accumulators acc[m,n] = sum_k a_q[m,k] * w_q[n,k] (the bit-exact hook).
    fp64 matmul is exact here: |acc| <= K * 127 * 128 < 2^53."""
    return (a_q.double() @ w_q.double().t()).to(torch.int64).to(torch.int32)


def int8_scaled_mm(a_q: torch.Tensor, w_q: torch.Tensor, scale_a: torch.Tensor, scale_b: torch.Tensor,
                   out_dtype: torch.dtype, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """native_w8a8_per_token_matmul (test/srt/cpu/utils.py:91-115) == torch_scaled_mm
    (sgl-kernel/tests/test_int8_gemm.py:13-19): C = As * (A @ B^T) * Bs (+ bias), fp32, then cast.
    a_q [M,K] int8; w_q [N,K] int8 (weight as stored by create_weights)."""
    acc = int8_accumulate(a_q, w_q).float()
    c = scale_a.view(-1, 1).float() * acc * scale_b.view(1, -1).float()
    if bias is not None:
        c = c + bias.view(1, -1).float()
    return c.to(out_dtype)


def w8a8_int8_linear(x: torch.Tensor, w_q: torch.Tensor, w_scale: torch.Tensor,
                     bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """W8A8Int8LinearMethod.apply (python/sglang/srt/layers/quantization/w8a8_int8.py:374-394)."""
    x2 = x.reshape(-1, x.shape[-1])
    x_q, x_s = per_token_quant_int8(x2)
    out = int8_scaled_mm(x_q, w_q, x_s, w_scale, x.dtype, bias)
    return out.reshape(x.shape[:-1] + (w_q.shape[0],))


# ---- block-scaled FP8 ----------------------------------------------------------------------------
def per_token_group_quant_fp8(x: torch.Tensor, group_size: int = 128, eps: float = 1e-10,
                              dtype: torch.dtype = torch.float8_e4m3fn) -> Tuple[torch.Tensor, torch.Tensor]:
    """native_per_token_group_quant_fp8 (python/sglang/test/test_block_fp8.py:24-49), with ONE rule fixed:
    the eps clamp is applied in fp32 as the reference's kernels do (fp8_kernel.py:145-150 loads y as fp32 and
    takes max(absmax, eps); per_token_group_quant_8bit.cu likewise).  The test restatement clamps in the input
    dtype, which only differs for all-zero groups (bf16 rounds eps to 1.0004e-10; fp16 flushes it to 0 and the
    restatement then produces NaN codes)."""
    assert x.shape[-1] % group_size == 0 and x.is_contiguous()
    finfo = torch.finfo(dtype)
    x_ = x.reshape(x.numel() // group_size, group_size)
    amax = x_.abs().max(dim=-1, keepdim=True)[0].to(torch.float32).clamp(min=eps)
    x_s = amax / finfo.max
    x_q = (x_ / x_s).clamp(min=finfo.min, max=finfo.max).to(dtype)
    return x_q.reshape(x.shape), x_s.reshape(x.shape[:-1] + (x.shape[-1] // group_size,))


def per_token_quant_fp8(x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """per_token_quant_fp8_kernel (sgl-kernel/csrc/gemm/per_token_quant_fp8.cu:15-90) restated the way
    its test does (sgl-kernel/tests/test_per_token_quant_fp8.py:14-22): scale = absmax/448,
    q = clamp(x * (1/scale))."""
    xf = x.float()
    s = xf.abs().amax(dim=-1, keepdim=True) / 448.0
    inv = torch.where(s == 0, torch.zeros_like(s), 1.0 / s)
    q = (xf * inv).clamp(-448.0, 448.0).to(torch.float8_e4m3fn)
    return q, s


def w8a8_block_fp8_matmul(a_q: torch.Tensor, w_q: torch.Tensor, a_s: torch.Tensor, w_s: torch.Tensor,
                          block_size=(128, 128), output_dtype: torch.dtype = torch.bfloat16) -> torch.Tensor:
    """native_w8a8_block_fp8_matmul (python/sglang/test/test_block_fp8.py:274-330), vectorised over
    the tile loops (same per-K-block fp32 partial products times As*Bs, summed in K-block order)."""
    block_n, block_k = block_size
    A = a_q.to(torch.float32).reshape(-1, a_q.shape[-1])
    B = w_q.to(torch.float32)
    M, K = A.shape
    N = B.shape[0]
    As = a_s.reshape(M, -1).float()
    k_tiles = (K + block_k - 1) // block_k
    C = torch.zeros((M, N), dtype=torch.float32)
    n_idx = torch.arange(N) // block_n
    for i in range(k_tiles):
        ks = slice(i * block_k, min((i + 1) * block_k, K))
        part = A[:, ks] @ B[:, ks].t()
        C += part * (As[:, i : i + 1] * w_s[n_idx, i].view(1, -1).float())
    return C.reshape(a_q.shape[:-1] + (N,)).to(output_dtype)


def w8a8_block_fp8_linear(x: torch.Tensor, w_q: torch.Tensor, w_s: torch.Tensor,
                          bias: Optional[torch.Tensor] = None, block_size=(128, 128)) -> torch.Tensor:
    """Fp8LinearMethod.apply block path (fp8.py:444-487 -> fp8_utils.py:180-210)."""
    x2 = x.reshape(-1, x.shape[-1]).contiguous()
    x_q, x_s = per_token_group_quant_fp8(x2, block_size[1])
    out = w8a8_block_fp8_matmul(x_q, w_q, x_s, w_s, block_size, x.dtype)
    if bias is not None:
        out = out + bias
    return out.to(x.dtype).reshape(x.shape[:-1] + (w_q.shape[0],))


def fp8_scaled_mm(a_q, w_q, scale_a, scale_b, out_dtype, bias=None):
    """Per-token x per-channel FP8 GEMM, restated as sgl-kernel/tests/test_fp8_gemm.py does
    (fp32 matmul * scale_a * scale_b + bias)."""
    o = a_q.float() @ w_q.float().t()
    o = o * scale_a.view(-1, 1).float() * scale_b.view(1, -1).float()
    if bias is not None:
        o = o + bias.float()
    return o.to(out_dtype)


# ---- AWQ int4 ------------------------------------------------------------------------------------------
def _reverse_awq_order(t: torch.Tensor) -> torch.Tensor:
    idx = torch.arange(t.shape[-1], dtype=torch.int32).view(-1, 8)[:, AWQ_REVERSE_ORDER].reshape(-1)
    return t[:, idx.long()] & 0xF


def awq_dequantize(qweight: torch.Tensor, scales: torch.Tensor, qzeros: torch.Tensor, group_size: int
                   ) -> torch.Tensor:
    """awq_dequantize_torch (sgl-kernel/tests/test_awq_dequant.py:27-61): nibble j of word c is logical
    column 8c + ORDER^-1..., W = (q - z) * s in the scale dtype."""
    if group_size == -1:
        group_size = qweight.shape[0]
    shifts = torch.arange(0, 32, 4)
    iw = torch.bitwise_right_shift(qweight[:, :, None], shifts[None, None, :]).to(torch.int8)
    iw = iw.view(iw.shape[0], -1)
    zs = torch.bitwise_right_shift(qzeros[:, :, None], shifts[None, None, :]).to(torch.int8)
    zs = zs.view(qzeros.shape[0], -1)
    zs = _reverse_awq_order(zs)
    iw = _reverse_awq_order(iw)
    iw = torch.bitwise_and(iw, 0xF)
    zs = torch.bitwise_and(zs, 0xF)
    sc = scales.repeat_interleave(group_size, dim=0)
    zs = zs.repeat_interleave(group_size, dim=0)
    return (iw - zs) * sc


def awq_linear(x: torch.Tensor, qweight: torch.Tensor, scales: torch.Tensor, qzeros: torch.Tensor,
               group_size: int, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """AWQLinearMethod.apply (python/sglang/srt/layers/quantization/awq.py:401-418): dequantise the
    whole matrix then torch.matmul; evaluated in fp32 on CPU and cast (fp16 matmul on CUDA accumulates
    in fp32 as well)."""
    w = awq_dequantize(qweight, scales, qzeros, group_size)
    x2 = x.reshape(-1, x.shape[-1])
    out = (x2.float() @ w.float())
    if bias is not None:
        out = out + bias.float()
    return out.to(x.dtype).reshape(x.shape[:-1] + (w.shape[1],))


# ---- element-wise steps either side of the path ---------------------------------------------------------
def rmsnorm(x: torch.Tensor, weight: torch.Tensor, eps: float, residual: Optional[torch.Tensor] = None):
    """RMSNorm.forward_native (python/sglang/srt/layers/layernorm.py:135-172)."""
    orig = x.dtype
    x = x.to(torch.float32)
    if residual is not None:
        x = x + residual.to(torch.float32)
        residual = x.to(orig)
    var = x.pow(2).mean(dim=-1, keepdim=True)
    x = x * torch.rsqrt(var + eps)
    x = (x * weight).to(orig)
    return x if residual is None else (x, residual)


def silu_and_mul(x: torch.Tensor) -> torch.Tensor:
    """SiluAndMul.forward_native (python/sglang/srt/layers/activation.py:60-62)."""
    d = x.shape[-1] // 2
    return torch.nn.functional.silu(x[..., :d]) * x[..., d:]


def rope_neox(positions: torch.Tensor, q: torch.Tensor, k: torch.Tensor, head_dim: int,
              cos_sin_cache: torch.Tensor):
    """RotaryEmbedding.forward_native with is_neox_style (rotary_embedding.py:139-167, :50-75), evaluated
    in fp32 with the fp32 cache as the CUDA path does (rotary_embedding.py:99-102)."""
    cos, sin = cos_sin_cache.index_select(0, positions).chunk(2, dim=-1)
    cos, sin = cos.unsqueeze(-2).float(), sin.unsqueeze(-2).float()

    def rot(x):
        shape = x.shape
        x = x.reshape(positions.shape[0], -1, head_dim).float()
        r = cos.shape[-1] * 2
        x1, x2 = x[..., : r // 2], x[..., r // 2 : r]
        o = torch.cat((x1 * cos - x2 * sin, x2 * cos + x1 * sin, x[..., r:]), dim=-1)
        return o.reshape(shape)

    return rot(q).to(q.dtype), rot(k).to(k.dtype)
