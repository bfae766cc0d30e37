"""Parity of the LSE-merge primitive and the element-wise steps either side of the path
against the oracle restatements; shape grids follow the reference's kernel tests
(sgl-kernel/tests/test_merge_state_v2.py, test_norm.py, test_activation.py, test_rotary_embedding.py)."""
import math

import pytest
import torch

from oracle import attention_oracle as ao
from oracle import quant_oracle as qo
from sglang_npu_b200 import kernels
from sglang_npu_b200.srt.model_executor.layer_stack import compute_cos_sin_cache
from tests.util import assert_close_path

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("T,H,D", [(1, 1, 64), (7, 14, 64), (64, 32, 128), (513, 40, 128), (3, 8, 256)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_merge_state_v2(cuda_device, T, H, D, dtype):
    g = torch.Generator().manual_seed(T * 131 + H)
    va, vb = torch.randn(T, H, D, generator=g).to(dtype), torch.randn(T, H, D, generator=g).to(dtype)
    sa, sb = torch.randn(T, H, generator=g) * 3, torch.randn(T, H, generator=g) * 3
    if T > 2:  # the reference test also feeds +-inf LSEs (empty partial states)
        sa[0, 0] = -math.inf
        sb[1, 0] = -math.inf
    ref_v, ref_s = ao.merge_state(va, sa, vb, sb)
    out_v, out_s = kernels.merge_state_v2(va.to(cuda_device), sa.to(cuda_device), vb.to(cuda_device), sb.to(cuda_device))
    assert_close_path(out_v, ref_v, "merge_state v")
    torch.testing.assert_close(out_s.cpu(), ref_s, rtol=1e-4, atol=1e-4)


def test_merge_state_is_associative_with_decode_splits(cuda_device):
    """merging the states of two halves of a KV range equals attention over the whole range (cascade.cu use)."""
    g = torch.Generator().manual_seed(5)
    T, H, D, N = 4, 8, 128, 96
    q = torch.randn(T, H, D, generator=g)
    k = torch.randn(T, N, H, D, generator=g)
    v = torch.randn(T, N, H, D, generator=g)

    def part(lo, hi):
        s = torch.einsum("thd,tnhd->thn", q, k[:, lo:hi]) * D ** -0.5
        lse = torch.logsumexp(s, dim=-1)
        return torch.einsum("thn,tnhd->thd", torch.softmax(s, -1), v[:, lo:hi]), lse

    (oa, sa), (ob, sb), (full, _) = part(0, 40), part(40, N), part(0, N)
    out, _ = kernels.merge_state_v2(oa.bfloat16().to(cuda_device), sa.to(cuda_device), ob.bfloat16().to(cuda_device),
                                    sb.to(cuda_device))
    assert_close_path(out, full.bfloat16(), "merge of halves")


@pytest.mark.parametrize("M,H", [(1, 896), (64, 4096), (37, 5120), (300, 8192), (5, 128)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_rmsnorm_and_fused_add(cuda_device, M, H, dtype):
    g = torch.Generator().manual_seed(M + H)
    x = torch.randn(M, H, generator=g).to(dtype)
    r = torch.randn(M, H, generator=g).to(dtype)
    w = (1 + 0.1 * torch.randn(H, generator=g)).to(dtype)
    out = kernels.rmsnorm(x.to(cuda_device), w.to(cuda_device), 1e-6)
    assert_close_path(out, qo.rmsnorm(x, w, 1e-6), "rmsnorm", max_abs=4e-2)
    xd, rd = x.to(cuda_device).clone(), r.to(cuda_device).clone()
    kernels.fused_add_rmsnorm(xd, rd, w.to(cuda_device), 1e-6)
    ref_x, ref_r = qo.rmsnorm(x, w, 1e-6, r)
    assert torch.equal(rd.cpu().view(torch.int16), ref_r.view(torch.int16)), "residual = bf16(x + residual) must be exact"
    assert_close_path(xd, ref_x, "fused_add_rmsnorm", max_abs=4e-2)


@pytest.mark.parametrize("M,H", [(1, 4864), (64, 14336), (129, 27648), (3, 64)])
def test_silu_and_mul(cuda_device, M, H):
    x = torch.randn(M, 2 * H, generator=torch.Generator().manual_seed(H)).bfloat16()
    out = kernels.silu_and_mul(x.to(cuda_device))
    # forward_native on the bf16 tensor: silu rounded to bf16, then the product rounded (activation.py:60-62);
    # one bf16 ulp (2^-8 relative) of slack for the fast-exp sigmoid
    torch.testing.assert_close(out.cpu().float(), qo.silu_and_mul(x).float(), rtol=1.6e-2, atol=2e-3)


@pytest.mark.parametrize("hq,hkv,d", [(32, 8, 128), (14, 2, 64), (5, 1, 128)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_rope_neox_on_strided_qkv(cuda_device, hq, hkv, d, dtype):
    T = 33
    g = torch.Generator().manual_seed(hq)
    qkv = torch.randn(T, (hq + 2 * hkv) * d, generator=g).to(dtype)
    pos = torch.randint(0, 4000, (T,), generator=g)
    cache = compute_cos_sin_cache(d, 4096, 1e6)
    q, k, v = qkv.split([hq * d, hkv * d, hkv * d], dim=-1)
    ref_q, ref_k = qo.rope_neox(pos, q, k, d, cache)
    dq = qkv.to(cuda_device)
    q2, k2, v2 = dq.split([hq * d, hkv * d, hkv * d], dim=-1)
    kernels.apply_rope_neox_inplace(pos.to(cuda_device), q2, k2, d, cache.to(cuda_device))
    assert_close_path(q2, ref_q, "rope q", max_abs=4e-2)
    assert_close_path(k2, ref_k, "rope k", max_abs=4e-2)
    assert torch.equal(v2.cpu(), v), "rope must not touch v"


# ---- fused steps: bit-identical to the two un-fused kernels run back to back ---------------------------------
@pytest.mark.parametrize("M,H", [(1, 896), (64, 4096), (37, 5120), (130, 8192)])
@pytest.mark.parametrize("with_residual", [False, True])
def test_rmsnorm_quant_int8_equals_unfused_and_oracle(cuda_device, M, H, with_residual):
    g = torch.Generator().manual_seed(M * 7 + H)
    x = torch.randn(M, H, generator=g).bfloat16().to(cuda_device)
    r = torch.randn(M, H, generator=g).bfloat16().to(cuda_device)
    w = (1 + 0.1 * torch.randn(H, generator=g)).bfloat16().to(cuda_device)
    # un-fused reference flow on the device
    x1, r1 = x.clone(), r.clone()
    if with_residual:
        kernels.fused_add_rmsnorm(x1, r1, w, 1e-5)
        y1 = x1
    else:
        y1 = kernels.rmsnorm(x1, w, 1e-5)
    q1, s1 = kernels.per_token_quant_int8(y1)
    r2 = r.clone()
    act, y2 = kernels.rmsnorm_quant_int8(x, w, 1e-5, residual=r2 if with_residual else None, return_normed=True)
    assert torch.equal(act.x_q, q1) and torch.equal(act.x_scale, s1) and torch.equal(y2, y1)
    if with_residual:
        assert torch.equal(r2, r1)
    # and against the oracle: norm (tolerance) -> quant of the device's own normed activation (bit-exact)
    qo_q, qo_s = qo.per_token_quant_int8(y2.cpu())
    assert torch.equal(act.x_q.cpu(), qo_q) and torch.equal(act.x_scale.cpu().view(-1), qo_s.view(-1))


@pytest.mark.parametrize("M,H", [(1, 4864), (64, 14336), (5, 27648), (300, 3584), (2, 64)])
def test_silu_and_mul_quant_int8_equals_unfused(cuda_device, M, H):
    x = torch.randn(M, 2 * H, generator=torch.Generator().manual_seed(H + M)).bfloat16().to(cuda_device)
    y = kernels.silu_and_mul(x)
    q1, s1 = kernels.per_token_quant_int8(y)
    act = kernels.silu_and_mul_quant_int8(x)
    assert torch.equal(act.x_q, q1) and torch.equal(act.x_scale, s1)


@pytest.mark.parametrize("hq,hkv,d", [(32, 8, 128), (14, 2, 64), (5, 1, 128)])
@pytest.mark.parametrize("loc_dtype", [torch.int64, torch.int32])
def test_rope_neox_set_kv_equals_unfused(cuda_device, hq, hkv, d, loc_dtype):
    T, slots = 41, 512
    g = torch.Generator().manual_seed(hq * 3 + d)
    qkv = torch.randn(T, (hq + 2 * hkv) * d, generator=g).bfloat16().to(cuda_device)
    pos = torch.randint(0, 4000, (T,), generator=g).to(cuda_device)
    loc = torch.randperm(slots - 16, generator=g)[:T].add(16).to(loc_dtype).to(cuda_device)
    cache = compute_cos_sin_cache(d, 4096, 5e5).to(cuda_device)
    a, b = qkv.clone(), qkv.clone()
    kc1 = torch.zeros(slots, hkv, d, dtype=torch.bfloat16, device=cuda_device)
    vc1, kc2, vc2 = torch.zeros_like(kc1), torch.zeros_like(kc1), torch.zeros_like(kc1)
    q1, k1, v1 = a.split([hq * d, hkv * d, hkv * d], dim=-1)
    kernels.apply_rope_neox_inplace(pos, q1, k1, d, cache)
    kernels.set_kv_buffer(kc1, vc1, loc, k1, v1)
    q2, k2, v2 = b.split([hq * d, hkv * d, hkv * d], dim=-1)
    kernels.rope_neox_set_kv(pos, q2, k2, v2, d, cache, kc2, vc2, loc)
    assert_close_path(q2, q1, "fused rope q", max_abs=1e-2)  # FMA contraction may differ by an ulp between kernels
    assert_close_path(k2, k1, "fused rope k", max_abs=1e-2)
    assert torch.equal(vc2, vc1), "v rows must be copied bit-exactly into the cache"
    assert torch.equal(kc2.view(slots, -1)[loc.long()], k2), "cache rows must equal the rotated k bit-exactly"
    touched = torch.zeros(slots, dtype=torch.bool, device=cuda_device)
    touched[loc.long()] = True
    assert not kc2[~touched].any() and not vc2[~touched].any()
    # oracle
    ref_q, ref_k = qo.rope_neox(pos.cpu(), qkv.cpu()[:, : hq * d], qkv.cpu()[:, hq * d: (hq + hkv) * d], d, cache.cpu())
    assert_close_path(q2, ref_q, "fused rope q vs oracle", max_abs=4e-2)
    assert_close_path(k2, ref_k, "fused rope k vs oracle", max_abs=4e-2)


def test_int8_linear_accepts_prequantised_activation(cuda_device):
    from sglang_npu_b200 import synth
    from sglang_npu_b200.srt.layers.linear import RowParallelLinear
    from sglang_npu_b200.srt.layers.quantization.w8a8_int8 import B200W8A8Int8Config

    with torch.device(cuda_device):
        lin = RowParallelLinear(1024, 512, bias=False, quant_config=B200W8A8Int8Config(), params_dtype=torch.bfloat16)
    w, s = synth.int8_weight(512, 1024, 3)
    lin.weight_loader(lin.weight, w.to(cuda_device))
    lin.weight_loader(lin.weight_scale, s.view(-1, 1).to(cuda_device))
    lin.quant_method.process_weights_after_loading(lin)
    x = torch.randn(9, 2048, generator=torch.Generator().manual_seed(0)).bfloat16().to(cuda_device)
    y_ref, _ = lin(kernels.silu_and_mul(x))
    y_fused, _ = lin(kernels.silu_and_mul_quant_int8(x))
    assert torch.equal(y_ref, y_fused)


@pytest.mark.parametrize("rows,V", [(1, 128256), (64, 128256), (7, 151936), (300, 2048), (3, 1000), (64, 16032)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_argmax_matches_torch(cuda_device, rows, V, dtype):
    g = torch.Generator().manual_seed(rows + V)
    x = torch.randn((rows, V), generator=g).to(dtype)
    x[0, V - 1] = 30.0  # maximum in the ragged tail
    if rows > 1:
        x[1, 5] = 31.0
        x[1, V // 2] = 31.0  # tie: the lowest index wins
    xd = x.to(cuda_device)
    for _ in range(2):  # second call checks the self-cleaning scratch
        out = kernels.argmax(xd)
    torch.cuda.synchronize()
    ref = torch.argmax(x.float(), dim=-1)
    if rows > 1:
        assert int(out[1]) == 5
        ref[1] = 5
    assert torch.equal(out.cpu(), ref)
    # strided rows (a [:, :V] slice of a padded matrix)
    pad = torch.full((rows, V + 8), 100.0, dtype=dtype, device=cuda_device)
    pad[:, :V] = xd
    assert torch.equal(kernels.argmax(pad[:, :V]).cpu(), ref)


# ---- decode prologue / epilogue pieces of SURVEY §8f #4: embedding gather, lm_head GEMM, vocab-parallel greedy ----------
@pytest.mark.parametrize("M,N,K", [(1, 1024, 512), (5, 4096, 896), (64, 16032, 4096), (200, 2056, 1024), (256, 1000, 5120)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_dense_gemm_nt_lm_head(cuda_device, M, N, K, dtype):
    """logits = hidden @ lm_head.T (logits_processor.py:470-500) on the weight-streaming tcgen05 kernel vs an fp32
    matmul of the same 16-bit operands; ragged N (not a multiple of the 128-row tile) and K (not of the 64-element
    TMA box) included.  Tolerance: one rounding of the fp32 sum to the output dtype."""
    g = torch.Generator().manual_seed(M * 7 + N)
    a = (torch.randn(M, K, generator=g) * 0.5).to(dtype)
    b = (torch.randn(N, K, generator=g) / math.sqrt(K)).to(dtype)
    ref = a.float() @ b.float().t()
    out = kernels.dense_gemm_nt(a.to(cuda_device), b.to(cuda_device))
    assert out.dtype == dtype and tuple(out.shape) == (M, N)
    ulp = 2.0 ** -8 if dtype == torch.bfloat16 else 2.0 ** -11
    tol = ref.abs() * ulp + 1e-3
    assert ((out.cpu().float() - ref).abs() <= tol).all(), (out.cpu().float() - ref).abs().max()
    assert_close_path(out, ref, "dense_gemm_nt")


def test_dense_gemm_nt_bias_and_bounds(cuda_device):
    g = torch.Generator().manual_seed(3)
    a = torch.randn(8, 256, generator=g).bfloat16()
    b = (torch.randn(384, 256, generator=g) / 16).bfloat16()
    bias = torch.randn(384, generator=g).bfloat16()
    out = kernels.dense_gemm_nt(a.to(cuda_device), b.to(cuda_device), bias=bias.to(cuda_device))
    assert_close_path(out, a.float() @ b.float().t() + bias.float(), "dense_gemm_nt + bias")
    with pytest.raises(RuntimeError, match="M 300 > 256"):
        kernels.dense_gemm_nt(torch.zeros(300, 256, dtype=torch.bfloat16, device=cuda_device), b.to(cuda_device))


@pytest.mark.parametrize("T,V,H", [(1, 100, 64), (64, 128256, 4096), (13, 5000, 896)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_embedding_gather_and_positions(cuda_device, T, V, H, dtype):
    """hidden = table[ids] bit-exact (a copy), positions = seq_lens - 1 in the same launch (forward_batch_info.py:960-962)."""
    g = torch.Generator(device=cuda_device).manual_seed(T)
    table = torch.randn(V, H, generator=g, device=cuda_device).to(dtype)
    ids = torch.randint(0, V, (T,), generator=g, device=cuda_device)
    ids[0] = V - 1
    seq = torch.randint(1, 4096, (T,), generator=g, device=cuda_device)
    pos = torch.full((T,), -7, dtype=torch.int64, device=cuda_device)
    out = kernels.embedding(ids, table, seq, pos)
    assert torch.equal(out, table[ids]) and torch.equal(pos, seq - 1)
    assert torch.equal(kernels.embedding(ids, table), table[ids])  # without the positions half


@pytest.mark.parametrize("tp", [2, 8])
@pytest.mark.parametrize("rows,V", [(1, 1000), (64, 128256), (7, 152064)])
def test_vocab_parallel_argmax_keys(cuda_device, rows, V, tp):
    """Per-rank keys max-combined == torch.argmax over the gathered logits (what logits_processor.py + sampler.py do),
    including exact ties across and inside ranks (lowest global index wins) and a ragged last rank."""
    g = torch.Generator(device=cuda_device).manual_seed(rows + tp)
    logits = torch.randn(rows, V, generator=g, device=cuda_device).bfloat16()
    if rows > 1:
        logits[1, V - 1] = 50.0     # the maximum on the last rank's last column
        logits[0, 5] = 60.0         # a tie between rank 0 ...
        logits[0, V - 3] = 60.0     # ... and the last rank
    per = (V + tp - 1) // tp // 8 * 8 + 8
    keys = None
    for r in range(tp):
        lo, hi = min(r * per, V), min((r + 1) * per, V)
        if hi <= lo:
            continue
        local = logits[:, lo:hi]
        assert local.stride(0) == V
        if (local.data_ptr() % 16) != 0:  # the kernel reads 16-byte vectors: a TP rank's slice is its own allocation
            local = local.contiguous()
        k = kernels.argmax_key(local, lo)
        assert (k >= 0).all()
        keys = k if keys is None else torch.maximum(keys, k)
    out = kernels.argmax_key_decode(keys)
    # expected: first index of the maximum (bf16 values are exact in fp32)
    exp = torch.argmax(logits.float(), dim=-1)
    mx = logits.float().amax(dim=-1, keepdim=True)
    first = torch.where(logits.float() == mx, torch.arange(V, device=cuda_device).expand(rows, V), V).amin(dim=-1)
    assert torch.equal(out, first), (out, first, exp)


@pytest.mark.parametrize("M,H", [(1, 4096), (64, 4096), (7, 5120), (300, 8192), (3, 128), (5, 4224)])
@pytest.mark.parametrize("with_residual", [False, True])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_rmsnorm_group_quant_fp8_equals_unfused(cuda_device, M, H, with_residual, dtype):
    """(fused_add_)rmsnorm -> per-(token, 128) e4m3 quant in one launch: same bits as the two kernels back to back
    (H = 4224 leaves the last pass with one active half-warp per warp)."""
    g = torch.Generator().manual_seed(M * 5 + H)
    x = torch.randn(M, H, generator=g).to(dtype).to(cuda_device)
    r = torch.randn(M, H, generator=g).to(dtype).to(cuda_device)
    w = (1 + 0.1 * torch.randn(H, generator=g)).to(dtype).to(cuda_device)
    if M > 2:
        x[2] = 0  # all-zero row: every group clamps at eps
        r[2] = 0
    x1, r1 = x.clone(), r.clone()
    if with_residual:
        kernels.fused_add_rmsnorm(x1, r1, w, 1e-5)
        y1 = x1
    else:
        y1 = kernels.rmsnorm(x1, w, 1e-5)
    q1, s1 = kernels.sgl_per_token_group_quant_fp8(y1, 128)
    r2 = r.clone()
    act = kernels.rmsnorm_group_quant_fp8(x, w, 1e-5, residual=r2 if with_residual else None)
    torch.cuda.synchronize()
    assert torch.equal(act.x_q.view(torch.uint8), q1.view(torch.uint8))
    assert torch.equal(act.x_scale.view(torch.int32), s1.view(torch.int32))
    if with_residual:
        assert torch.equal(r2, r1)


@pytest.mark.parametrize("M,H", [(1, 14336), (64, 14336), (5, 27648), (300, 3584), (2, 128), (8192, 1792)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_silu_and_mul_group_quant_fp8_equals_unfused(cuda_device, M, H, dtype):
    x = torch.randn(M, 2 * H, generator=torch.Generator().manual_seed(H + M)).to(dtype).to(cuda_device)
    y = kernels.silu_and_mul(x)
    q1, s1 = kernels.sgl_per_token_group_quant_fp8(y, 128)
    act = kernels.silu_and_mul_group_quant_fp8(x)
    torch.cuda.synchronize()
    assert torch.equal(act.x_q.view(torch.uint8), q1.view(torch.uint8))
    assert torch.equal(act.x_scale.view(torch.int32), s1.view(torch.int32))
