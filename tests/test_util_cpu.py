"""The parity assertion itself: it must catch a dropped KV block on a long-context row (a global cosine
plus an absolute 2e-2 bound let one through)."""
import pytest
import torch

from oracle import attention_oracle as ao
from sglang_npu_b200 import synth
from tests.util import assert_close_path


def _decode(lens, k, v, pb, q):
    return ao.decode_attention(q, k, v, pb.req_to_token, pb.req_pool_indices, pb.seq_lens, q.shape[-1] ** -0.5,
                               compute_dtype=torch.float32)


def test_assert_fires_when_one_kv_block_of_the_longest_request_is_perturbed():
    lens = [1, 16, 255, 700, 1030]
    hq, hkv, d, page = 8, 2, 64, 16
    pb = synth.make_paged_batch(lens, page, seed=5)
    k, v = synth.randn_kv(pb.num_slots, page, hkv, d, torch.bfloat16, seed=6)
    q = torch.randn(len(lens), hq, d, generator=torch.Generator().manual_seed(7)).to(torch.bfloat16)
    ref = _decode(lens, k, v, pb, q)
    assert_close_path(ref.to(torch.bfloat16), ref, "bf16 rounding of the exact result passes")
    # drop one 16-token block of the 1030-token request (what a mis-merged split or a skipped page looks like)
    slots = pb.req_to_token[pb.req_pool_indices[4], 512:528].long()
    v2 = v.clone()
    v2[slots] = 0
    bad = _decode(lens, k, v2, pb, q)
    err = (bad - ref).abs().max().item()
    assert err < 2e-2, "the perturbation must sit inside the absolute tolerance for this test to mean anything"
    with pytest.raises(AssertionError, match="relative error|per-row cosine"):
        assert_close_path(bad, ref, "one dropped block")


def test_assert_accepts_near_zero_rows_on_the_global_scale():
    a = torch.zeros(4, 8)
    a[0] = 1.0
    b = a.clone()
    b[1] += 1e-5  # an all-zero expected row with noise far below the global scale
    assert_close_path(b, a, "zero rows")
