"""Golden vectors for the page-table kernels, produced by the UNMODIFIED reference (this container only).

    python tests/golden/make_golden_page_table.py

Runs the reference's own torch implementations — alloc_extend_native / alloc_decode_kernel_ascend
(python/sglang/srt/mem_cache/allocator.py:573-660, the restatements the reference ships next to its Triton kernels)
and compute_position_torch (model_executor/forward_batch_info.py:942-957) — imported from /root/reference through
oracle/ref_shims.py, on seeded allocator states, and stores inputs + outputs in tests/golden/page_table_*.pt.
"""
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import ref_shims  # noqa: E402


def make_state(g, bs, page, max_len, decode):
    """A consistent allocator state: every request owns pages for its prefix; last_loc is the slot of its last
    prefix token; free_pages is a shuffled list of the remaining pages."""
    num_pages = bs * (2 * max_len // page + 3) + 8
    perm = (torch.randperm(num_pages, generator=g) + 1).tolist()
    pre_lens, seq_lens, last_loc = [], [], []
    for b in range(bs):
        kind = int(torch.randint(0, 5, (1,), generator=g))
        pre = int(torch.randint(0, max_len, (1,), generator=g))
        if kind == 0:
            pre = 0
        elif kind == 1:
            pre = (pre // page) * page  # on a page boundary
        ext = 1 if decode else int(torch.randint(1, max_len, (1,), generator=g))
        if not decode and kind == 2:
            ext = max(1, min(ext, (page - pre % page) % page))  # stays inside the old partial page
        n_pages = -(-pre // page)
        pages = [perm.pop() for _ in range(n_pages)]
        pre_lens.append(pre)
        seq_lens.append(pre + ext)
        last_loc.append(pages[-1] * page + (pre - 1) % page if pre > 0 else -1)
    free_pages = torch.tensor(perm, dtype=torch.int64)
    return (torch.tensor(pre_lens, dtype=torch.int64), torch.tensor(seq_lens, dtype=torch.int64),
            torch.tensor(last_loc, dtype=torch.int64), free_pages)


def main():
    ref_shims.install()
    from sglang.srt.mem_cache.allocator import alloc_decode_kernel_ascend, alloc_extend_native
    from sglang.srt.model_executor.forward_batch_info import compute_position_torch

    g = torch.Generator().manual_seed(1234)
    cases = []
    for page in (1, 16, 64):
        for bs in (1, 3, 8, 33):
            pre, seq, ll, free = make_state(g, bs, page, 200, decode=False)
            n_new = int(((seq + page - 1) // page - (pre + page - 1) // page).sum())
            free = free[: n_new + 4].clone()
            out, _ = alloc_extend_native(pre, seq, ll, free, page, n_new)
            pos, start = compute_position_torch(pre.to(torch.int32), (seq - pre).to(torch.int32))
            cases.append(dict(kind="extend", page=page, prefix_lens=pre, seq_lens=seq, last_loc=ll, free_pages=free,
                              out_indices=out.to(torch.int64), new_pages=n_new, positions=pos,
                              extend_start_loc=start.to(torch.int32)))
            pre, seq, ll, free = make_state(g, bs, page, 200, decode=True)
            npp = (seq + page - 1) // page - (seq - 1 + page - 1) // page
            free = free[: int(npp.sum()) + 4].clone()
            out = torch.empty(bs, dtype=torch.int64)
            alloc_decode_kernel_ascend(seq, ll, free, out, page, npp)
            cases.append(dict(kind="decode", page=page, seq_lens=seq, last_loc=ll, free_pages=free, out_indices=out,
                              new_pages=int(npp.sum())))
    path = os.path.join(HERE, "page_table_alloc.pt")
    torch.save(cases, path)
    print(f"wrote {path}: {len(cases)} cases, {os.path.getsize(path)} bytes")


if __name__ == "__main__":
    main()
