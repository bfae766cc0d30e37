"""GPU parity (bit-exact) of the page-table kernels (csrc/page_table.cu) against the reference-generated goldens and
the oracle, the fused last_loc / write-back variants, and the allocator class end to end."""
import os

import pytest
import torch

from oracle import page_table_oracle as po
from sglang_npu_b200 import kernels
from sglang_npu_b200.srt.mem_cache.allocator import PagedTokenToKVPoolAllocator
from tests.util import GOLDEN_DIR

pytestmark = pytest.mark.gpu

CASES = torch.load(os.path.join(GOLDEN_DIR, "page_table_alloc.pt"), weights_only=False)


@pytest.mark.parametrize("lens_dtype", [torch.int64, torch.int32])
def test_alloc_extend_goldens(cuda_device, lens_dtype):
    dev = cuda_device
    for c in [c for c in CASES if c["kind"] == "extend"]:
        out = torch.full((c["out_indices"].numel(),), -7, dtype=torch.int64, device=dev)
        ret = torch.zeros((), dtype=torch.int64, device=dev)
        kernels.alloc_extend(c["prefix_lens"].to(dev, lens_dtype), c["seq_lens"].to(dev, lens_dtype),
                             c["last_loc"].to(dev, lens_dtype), c["free_pages"].to(dev), out, ret, c["page"])
        assert torch.equal(out.cpu(), c["out_indices"])
        r = int(ret.item())
        assert r >> 32 == c["new_pages"] and (r & 0xFFFFFFFF) == c["out_indices"].numel()


def test_alloc_decode_goldens(cuda_device):
    dev = cuda_device
    for c in [c for c in CASES if c["kind"] == "decode"]:
        out = torch.full((c["out_indices"].numel(),), -7, dtype=torch.int64, device=dev)
        ret = torch.zeros((), dtype=torch.int64, device=dev)
        kernels.alloc_decode(c["seq_lens"].to(dev), c["last_loc"].to(dev), c["free_pages"].to(dev), out, ret, c["page"])
        assert torch.equal(out.cpu(), c["out_indices"])
        assert int(ret.item()) == c["new_pages"]


def test_compute_position_goldens(cuda_device):
    dev = cuda_device
    for c in [c for c in CASES if c["kind"] == "extend"]:
        pre, ext = c["prefix_lens"].to(dev, torch.int32), (c["seq_lens"] - c["prefix_lens"]).to(dev, torch.int32)
        pos, start = kernels.compute_position(pre, ext, int(ext.sum()))
        assert torch.equal(pos.cpu(), c["positions"]) and torch.equal(start.cpu(), c["extend_start_loc"])
        pos0, _ = kernels.compute_position(pre[:0], ext, int(ext.sum()))  # no prefix tensor -> prefix 0
        assert torch.equal(pos0.cpu(), po.compute_position(torch.zeros_like(ext).cpu(), ext.cpu())[0])


def _state(seed, bs, page, width=512):
    g = torch.Generator().manual_seed(seed)
    r2t = torch.zeros((bs + 3, width), dtype=torch.int32)
    pool = torch.randperm(bs + 3, generator=g)[:bs]
    pre = torch.randint(0, 200, (bs,), generator=g)
    pre[0] = 0
    ext = torch.randint(1, 200, (bs,), generator=g)
    num_pages = int(((pre + ext + page - 1) // page).sum()) + 8
    perm = (torch.randperm(num_pages, generator=g) + 1).tolist()
    for r, p in zip(pool.tolist(), pre.tolist()):
        for t in range(p):
            if t % page == 0:
                pg = perm.pop()
            r2t[r, t] = pg * page + t % page
    return r2t, pool, pre, pre + ext, torch.tensor(perm, dtype=torch.int64)


@pytest.mark.parametrize("page", [1, 16, 64])
@pytest.mark.parametrize("bs", [1, 7, 64, 300])
def test_fused_extend_equals_three_step_path(cuda_device, page, bs):
    """last_loc read in-kernel + write-back == get_last_loc -> alloc_extend -> write_req_to_token_pool (oracle)."""
    dev = cuda_device
    r2t, pool, pre, seq, free = _state(page * 1000 + bs, bs, page)
    last = po.get_last_loc(r2t, pool, pre)
    ref_out, ref_pages, total = po.alloc_extend(pre, seq, last, free, page)
    ref_r2t = r2t.clone()
    po.write_req_to_token_pool(ref_r2t, pool, pre, seq, ref_out)
    # separate kernels
    d_r2t = r2t.to(dev)
    d_last = kernels.get_last_loc(d_r2t, pool.to(dev), pre.to(dev))
    assert torch.equal(d_last.cpu(), last)
    out = torch.empty(total, dtype=torch.int64, device=dev)
    ret = torch.zeros((), dtype=torch.int64, device=dev)
    kernels.alloc_extend(pre.to(dev), seq.to(dev), d_last, free.to(dev), out, ret, page)
    assert torch.equal(out.cpu(), ref_out)
    kernels.write_req_to_token_pool(d_r2t, pool.to(dev), pre.to(dev), seq.to(dev), (seq - pre).to(dev), out)
    assert torch.equal(d_r2t.cpu(), ref_r2t)
    # one fused launch
    f_r2t = r2t.to(dev)
    out2 = torch.empty(total, dtype=torch.int64, device=dev)
    kernels.alloc_extend(pre.to(dev), seq.to(dev), None, free.to(dev), out2, ret, page, req_to_token=f_r2t,
                         req_pool_indices=pool.to(dev), write_req_to_token=True)
    assert torch.equal(out2.cpu(), ref_out) and torch.equal(f_r2t.cpu(), ref_r2t)
    assert int(ret.item()) >> 32 == ref_pages
    # CSR view of the finished table (create_flashinfer_kv_indices)
    indptr = torch.zeros(bs + 1, dtype=torch.int32)
    indptr[1:] = torch.cumsum(seq, 0)
    kv = torch.empty(int(seq.sum()), dtype=torch.int32, device=dev)
    kernels.create_flashinfer_kv_indices(f_r2t, pool.to(dev), seq.to(dev), indptr.to(dev), None, kv)
    assert torch.equal(kv.cpu(), po.create_kv_indices(ref_r2t, pool, seq, indptr, None, int(seq.sum())))
    start = torch.minimum(pre, torch.full_like(pre, 3)).to(torch.int32)
    lens2 = (seq - start).to(torch.int32)
    indptr2 = torch.zeros(bs + 1, dtype=torch.int32)
    indptr2[1:] = torch.cumsum(lens2, 0)
    kv2 = torch.empty(int(lens2.sum()), dtype=torch.int32, device=dev)
    kernels.create_flashinfer_kv_indices(f_r2t, pool.to(dev, torch.int32), lens2.to(dev), indptr2.to(dev), start.to(dev), kv2)
    assert torch.equal(kv2.cpu(), po.create_kv_indices(ref_r2t, pool, lens2, indptr2, start, int(lens2.sum())))


@pytest.mark.parametrize("page", [1, 16])
def test_allocator_decode_loop_matches_oracle(cuda_device, page):
    """Prefill three requests, then 40 fused decode steps: req_to_token stays identical to the oracle's."""
    dev = cuda_device
    alloc = PagedTokenToKVPoolAllocator(4096, page, torch.bfloat16, dev)
    free_ref = alloc.free_pages.cpu().clone()
    r2t = torch.zeros((4, 256), dtype=torch.int32, device=dev)
    ref_r2t = r2t.cpu().clone()
    pool = torch.tensor([2, 0, 3], dtype=torch.int64)
    pre = torch.zeros(3, dtype=torch.int64)
    seq = torch.tensor([17, 1, 48], dtype=torch.int64)
    out = alloc.alloc_extend_fused(r2t, pool.to(dev), pre.to(dev), seq.to(dev), int(seq.sum()))
    ref_out, n_pages, _ = po.alloc_extend(pre, seq, po.get_last_loc(ref_r2t, pool, pre), free_ref, page)
    po.write_req_to_token_pool(ref_r2t, pool, pre, seq, ref_out)
    free_ref = free_ref[n_pages:]
    assert torch.equal(out.cpu(), ref_out)
    for _ in range(40):
        seq = seq + 1
        out = alloc.alloc_decode_fused(r2t, pool.to(dev), seq.to(dev))
        ref_out, n_pages = po.alloc_decode(seq, po.get_last_loc(ref_r2t, pool, seq - 1), free_ref, page)
        po.write_req_to_token_pool(ref_r2t, pool, seq - 1, seq, ref_out)
        free_ref = free_ref[n_pages:]
        assert torch.equal(out.cpu(), ref_out)
    assert torch.equal(r2t.cpu(), ref_r2t)
    assert torch.equal(alloc.free_pages.cpu(), free_ref)
    # freeing returns whole pages and they come back sorted
    alloc.free(out)
    before = alloc.available_size()
    alloc.merge_and_sort_free()
    assert alloc.available_size() == before and bool((alloc.free_pages[1:] > alloc.free_pages[:-1]).all())
    assert alloc.alloc(alloc.available_size() + page) is None
