"""Full-size parity on BASELINE config 3's shapes (Llama-3.1-8B, 8192-token prefill chunk, 6144-token cached prefix).

The oracle cannot evaluate an 8192 x 28672 x 4096 GEMM or an 8k causal attention in seconds, so each case checks
(a) SAMPLED rows against the oracle on exactly the operands the kernel saw (int32 accumulators bit-exact, outputs
within the north_star tolerance), and (b) a size-independent property over the WHOLE output (row-permutation
equivariance for the GEMMs — any mis-walked tile of the persistent schedule moves rows; chunk-invariance and
determinism for attention)."""
import os

import pytest
import torch

from oracle import quant_oracle as qo
from sglang_npu_b200 import kernels, synth
from tests import harness
from tests.util import assert_close_path

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(os.environ.get("B200_SKIP_FULLSIZE") == "1", reason="full-size cases skipped by env")]

LLAMA8B_KN = [(4096, 6144), (4096, 4096), (4096, 28672), (14336, 4096)]  # qkv, o, gate_up, down  [K, N]
M_FULL = 8192


def _sample_rows(M, n=24, seed=0):
    g = torch.Generator().manual_seed(seed)
    # tile corners of the 128-row MMA tiles + random interior rows
    fixed = [0, 1, 127, 128, 129, M // 2 - 1, M // 2, M - 129, M - 128, M - 1]
    rnd = torch.randint(0, M, (n - len(fixed),), generator=g).tolist()
    return torch.tensor(sorted(set(fixed + rnd)), dtype=torch.long)


@pytest.mark.parametrize("K,N", LLAMA8B_KN)
def test_int8_gemm_m8192(cuda_device, K, N):
    dev = cuda_device
    g = torch.Generator(device=dev).manual_seed(K + N)
    x = (torch.randn(M_FULL, K, generator=g, device=dev) / 10).to(torch.bfloat16)
    w = torch.randint(-127, 128, (N, K), generator=g, device=dev, dtype=torch.int32).to(torch.int8)
    ws = torch.rand(N, generator=g, device=dev) * 1e-2
    xq, xs = kernels.per_token_quant_int8(x)
    out, acc = kernels.int8_scaled_mm(xq, w.t(), xs, ws, torch.bfloat16, return_acc=True)
    rows = _sample_rows(M_FULL, seed=N)
    # (a) sampled rows: quant codes + int32 accumulators bit-exact, outputs vs the oracle
    rq, rs = qo.per_token_quant_int8(x[rows].cpu())
    assert torch.equal(xq[rows].cpu(), rq) and torch.equal(xs[rows].cpu(), rs)
    w_cpu = w.cpu()
    exact = qo.int8_accumulate(rq, w_cpu)
    assert torch.equal(acc[rows].cpu(), exact)
    ref = qo.int8_scaled_mm(rq, w_cpu, rs, ws.cpu(), torch.float32)
    peak = ref.abs().max().item()
    assert_close_path(out[rows], ref, f"int8 M=8192 K={K} N={N}", max_abs=max(2e-2, peak * 2.0 ** -8))
    # (b) whole output: a row permutation of the activations permutes the output rows, bit for bit
    perm = torch.randperm(M_FULL, generator=g, device=dev)
    out_p = kernels.int8_scaled_mm(xq[perm].contiguous(), w.t(), xs[perm].contiguous(), ws, torch.bfloat16)
    assert torch.equal(out_p, out[perm])
    del acc


@pytest.mark.parametrize("K,N", LLAMA8B_KN)
def test_block_fp8_gemm_m8192(cuda_device, K, N):
    dev = cuda_device
    g = torch.Generator(device=dev).manual_seed(K * 3 + N)
    x = torch.randn(M_FULL, K, generator=g, device=dev).to(torch.bfloat16)
    w = ((torch.rand(N, K, generator=g, device=dev) - 0.5) * 2 * 448).to(torch.float8_e4m3fn)
    sb = torch.rand(N // 128, K // 128, generator=g, device=dev) * 1e-3 + 1e-4
    xq, xs = kernels.sgl_per_token_group_quant_fp8(x, 128)
    out = kernels.fp8_blockwise_scaled_mm(xq, w.t(), xs, sb.t(), torch.bfloat16)
    rows = _sample_rows(M_FULL, seed=K)
    rq, rs = qo.per_token_group_quant_fp8(x[rows].cpu().contiguous(), 128)
    assert torch.equal(xq[rows].cpu().view(torch.uint8), rq.view(torch.uint8)) and torch.equal(xs[rows].cpu(), rs)
    ref = qo.w8a8_block_fp8_matmul(rq, w.cpu(), rs, sb.cpu(), (128, 128), torch.float32)
    peak = ref.abs().max().item()
    assert_close_path(out[rows], ref, f"block fp8 M=8192 K={K} N={N}", max_abs=max(2e-2, peak * 2.0 ** -7))
    perm = torch.randperm(M_FULL, generator=g, device=dev)
    out_p = kernels.fp8_blockwise_scaled_mm(xq[perm].contiguous(), w.t(), xs[perm].contiguous(), sb.t(), torch.bfloat16)
    assert torch.equal(out_p, out[perm])


@pytest.mark.parametrize("K,N", LLAMA8B_KN)
def test_channel_fp8_gemm_m8192(cuda_device, K, N):
    dev = cuda_device
    g = torch.Generator(device=dev).manual_seed(K * 5 + N)
    x = torch.randn(M_FULL, K, generator=g, device=dev).to(torch.bfloat16)
    w = ((torch.rand(N, K, generator=g, device=dev) - 0.5) * 2 * 448).to(torch.float8_e4m3fn)
    sb = torch.rand(N, generator=g, device=dev) * 1e-4
    xq, xs = kernels.sgl_per_token_quant_fp8(x)
    out = kernels.fp8_scaled_mm(xq, w.t(), xs, sb, torch.bfloat16)
    rows = _sample_rows(M_FULL, seed=K + 1)
    ref = qo.fp8_scaled_mm(xq[rows].cpu(), w.cpu(), xs[rows].cpu(), sb.cpu(), torch.float32)
    peak = ref.abs().max().item()
    assert_close_path(out[rows], ref, f"per-channel fp8 M=8192 K={K} N={N}", max_abs=max(2e-2, peak * 2.0 ** -7))
    perm = torch.randperm(M_FULL, generator=g, device=dev)
    out_p = kernels.fp8_scaled_mm(xq[perm].contiguous(), w.t(), xs[perm].contiguous(), sb, torch.bfloat16)
    assert torch.equal(out_p, out[perm])


def _attn_rows_oracle(q_rows, pos_rows, k_all, v_all, hq, hkv, d):
    """softmax(q k^T / sqrt(d)) v for single query rows: row i sees keys [0, pos_rows[i]] of its request (the
    reference formula, torch_native_backend.py:27-110, evaluated in fp32 for the sampled rows only)."""
    grp = hq // hkv
    out = torch.empty(q_rows.shape[0], hq, d)
    for i, p in enumerate(pos_rows.tolist()):
        k = k_all[: p + 1].float()  # [p+1, hkv, d]
        v = v_all[: p + 1].float()
        qi = q_rows[i].float().view(hkv, grp, d)
        s = torch.einsum("hgd,thd->hgt", qi, k) * d ** -0.5
        pr = torch.softmax(s, dim=-1)
        out[i] = torch.einsum("hgt,thd->hgd", pr, v).reshape(hq, d)
    return out


@pytest.mark.parametrize("prefix,new", [(0, 8192), (6144, 2048)])
def test_extend_attention_cfg3(cuda_device, prefix, new):
    """8192-token causal prefill and the chunked case with 6144 cached (radix-matched) tokens, Llama-8B heads, page 16."""
    dev = cuda_device
    hq, hkv, d, page = 32, 8, 128, 16
    total = prefix + new
    pb = synth.make_paged_batch([total], page, seed=31)
    runner, backend = harness.build_runner(dev, pb.req_to_token, pb.num_slots, page, hq, hkv, d, torch.bfloat16, ctx=total)
    pool = runner.token_to_kv_pool
    kc, vc = synth.randn_kv_device(pb.num_slots, page, hkv, d, torch.bfloat16, dev, seed=32)
    pool.k_buffer[0].copy_(kc)
    pool.v_buffer[0].copy_(vc)
    g = torch.Generator(device=dev).manual_seed(33)
    q = torch.randn(new, hq * d, generator=g, device=dev).to(torch.bfloat16)
    k = torch.randn(new, hkv * d, generator=g, device=dev).to(torch.bfloat16)
    v = torch.randn(new, hkv * d, generator=g, device=dev).to(torch.bfloat16)
    slots = pb.req_to_token[pb.req_pool_indices[0], :total].long()
    fb = harness.extend_batch(runner, backend, pb.req_pool_indices, torch.tensor([total]), torch.tensor([prefix]),
                              slots[prefix:])
    layer = harness.make_layer(hq, hkv, d)
    o = layer(q, k, v, fb).view(new, hq, d)
    torch.cuda.synchronize()
    # KV write bit-exact over the whole chunk
    assert torch.equal(pool.k_buffer[0][slots[prefix:].to(dev)].view(new, -1), k)
    assert torch.equal(pool.v_buffer[0][slots[prefix:].to(dev)].view(new, -1), v)
    # (a) sampled query rows (tile corners + random) vs the oracle formula on the post-write cache
    rows = _sample_rows(new, n=20, seed=prefix)
    k_all = pool.k_buffer[0][slots.to(dev)].cpu()
    v_all = pool.v_buffer[0][slots.to(dev)].cpu()
    ref = _attn_rows_oracle(q[rows].cpu(), rows + prefix, k_all, v_all, hq, hkv, d)
    assert_close_path(o[rows], ref, f"extend cfg3 prefix={prefix} new={new}")
    # (b) determinism + chunk invariance over ALL rows: the same tokens extended as two halves (second half sees the
    # first through the cache) agree with the single launch
    o_again = layer(q, k, v, fb, save_kv_cache=False, kv_in_cache=True).view(new, hq, d)
    assert torch.equal(o, o_again)
    # the same chunk with its K/V read from the activations instead of the cache: the one-tile cp.async kernel
    o_act = layer(q, k, v, fb, save_kv_cache=False).view(new, hq, d)
    assert_close_path(o_act, o.float(), "new tokens from the activations vs from the cache")
    half = new // 2
    fb1 = harness.extend_batch(runner, backend, pb.req_pool_indices, torch.tensor([prefix + half]), torch.tensor([prefix]),
                               slots[prefix: prefix + half])
    o1 = layer(q[:half], k[:half], v[:half], fb1, save_kv_cache=False).view(half, hq, d)
    fb2 = harness.extend_batch(runner, backend, pb.req_pool_indices, torch.tensor([total]), torch.tensor([prefix + half]),
                               slots[prefix + half:])
    o2 = layer(q[half:], k[half:], v[half:], fb2, save_kv_cache=False).view(new - half, hq, d)
    torch.cuda.synchronize()
    assert_close_path(torch.cat([o1, o2]), o, "two chunks vs one launch", max_abs=1e-2)
