"""CPU: the page-table oracle (oracle/page_table_oracle.py) against golden vectors produced by the reference's own
torch implementations (tests/golden/make_golden_page_table.py), plus properties the allocator must keep."""
import os

import torch

from oracle import page_table_oracle as po
from tests.util import GOLDEN_DIR

CASES = torch.load(os.path.join(GOLDEN_DIR, "page_table_alloc.pt"), weights_only=False)


def test_alloc_extend_matches_reference_goldens():
    n = 0
    for c in CASES:
        if c["kind"] != "extend":
            continue
        out, new_pages, total = po.alloc_extend(c["prefix_lens"], c["seq_lens"], c["last_loc"], c["free_pages"], c["page"])
        assert torch.equal(out, c["out_indices"]), f"page {c['page']} bs {len(c['seq_lens'])}"
        assert new_pages == c["new_pages"] and total == int((c["seq_lens"] - c["prefix_lens"]).sum())
        n += 1
    assert n == 12


def test_alloc_decode_matches_reference_goldens():
    n = 0
    for c in CASES:
        if c["kind"] != "decode":
            continue
        out, new_pages = po.alloc_decode(c["seq_lens"], c["last_loc"], c["free_pages"], c["page"])
        assert torch.equal(out, c["out_indices"])
        assert new_pages == c["new_pages"]
        n += 1
    assert n == 12


def test_compute_position_matches_reference_goldens():
    for c in CASES:
        if c["kind"] != "extend":
            continue
        pre, ext = c["prefix_lens"].to(torch.int32), (c["seq_lens"] - c["prefix_lens"]).to(torch.int32)
        pos, start = po.compute_position(pre, ext)
        assert torch.equal(pos, c["positions"]) and torch.equal(start, c["extend_start_loc"])


def test_alloc_extend_properties():
    """Slots are unique, page-consistent with the request's last prefix slot, and contiguous inside a page."""
    for c in CASES:
        if c["kind"] != "extend":
            continue
        P = c["page"]
        out, _, _ = po.alloc_extend(c["prefix_lens"], c["seq_lens"], c["last_loc"], c["free_pages"], P)
        assert len(torch.unique(out)) == len(out)
        pt = 0
        for pre, seq, ll in zip(c["prefix_lens"].tolist(), c["seq_lens"].tolist(), c["last_loc"].tolist()):
            slots = out[pt:pt + seq - pre].tolist()
            pt += seq - pre
            for j, s in enumerate(slots):
                assert s % P == (pre + j) % P  # token position inside its page
                if j and (pre + j) % P:
                    assert s == slots[j - 1] + 1
            if pre % P:
                assert slots[0] == ll + 1


def test_get_last_loc_write_back_and_kv_indices_roundtrip():
    g = torch.Generator().manual_seed(3)
    r2t = torch.zeros((6, 64), dtype=torch.int32)
    pool = torch.tensor([4, 1, 5], dtype=torch.int64)
    pre = torch.tensor([0, 7, 20], dtype=torch.int64)
    seq = torch.tensor([9, 8, 41], dtype=torch.int64)
    for r, p in zip(pool.tolist(), pre.tolist()):
        r2t[r, :p] = torch.randint(1, 1000, (p,), generator=g, dtype=torch.int32)
    loc = torch.randint(1000, 2000, (int((seq - pre).sum()),), generator=g)
    last = po.get_last_loc(r2t, pool, pre)
    assert last.tolist() == [-1, int(r2t[1, 6]), int(r2t[5, 19])]
    po.write_req_to_token_pool(r2t, pool, pre, seq, loc)
    indptr = torch.tensor([0, 9, 17], dtype=torch.int32)
    kv = po.create_kv_indices(r2t, pool, seq, indptr, None, int(seq.sum()))
    assert kv[:9].tolist() == loc[:9].tolist()  # request 0 had no prefix: all its slots are the new ones
    assert kv[9 + 7] == loc[9] and kv[17 + 20:].tolist() == loc[10:].tolist()
