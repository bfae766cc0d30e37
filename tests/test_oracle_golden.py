"""CPU: pin the oracle restatement against golden vectors produced by the REAL reference
(tests/golden/make_golden.py ran TorchNativeAttnBackend et al. from /root/reference)."""
import pytest
import torch

from oracle import attention_oracle as ao
from tests.util import golden_files, load_golden

FILES = golden_files("attn_")


def test_golden_files_present():
    assert len(FILES) >= 5


@pytest.mark.parametrize("path", FILES, ids=lambda p: p.split("attn_")[-1][:-3])
def test_oracle_matches_reference_golden(path):
    rec = load_golden(path)
    c = rec["case"]
    k_cache = rec["k_cache_before"].clone()
    v_cache = rec["v_cache_before"].clone()
    k = rec["k"].view(-1, c["hkv"], c["d"])
    v = rec["v"].view(-1, c["hkv"], c["d"])
    # KV write: bit-exact against the reference's MHATokenToKVPool.set_kv_buffer
    ao.set_kv_buffer(k_cache, v_cache, rec["out_cache_loc"], k, v)
    assert torch.equal(k_cache.view(torch.int16), rec["k_cache_after"].view(torch.int16))
    assert torch.equal(v_cache.view(torch.int16), rec["v_cache_after"].view(torch.int16))
    q = rec["q"].view(-1, c["hq"], c["d"])
    scaling = c["d"] ** -0.5
    if c["mode"] == "decode":
        o = ao.decode_attention(q, k_cache, v_cache, rec["req_to_token"], rec["req_pool_indices"], rec["seq_lens"],
                                scaling)
    else:
        o = ao.extend_attention(q, k_cache, v_cache, rec["req_to_token"], rec["req_pool_indices"], rec["seq_lens"],
                                rec["extend_prefix_lens"], rec["extend_seq_lens"], scaling)
    # same torch SDPA call sequence on the same machine -> identical bits
    assert torch.equal(o.reshape(rec["o"].shape).view(torch.int16), rec["o"].view(torch.int16))


@pytest.mark.parametrize("path", FILES[:2], ids=lambda p: p.split("attn_")[-1][:-3])
def test_fp32_oracle_close_to_reference(path):
    """The fp32 evaluation of the same formula (used as the tight yardstick on GPU) agrees with the
    reference's bf16 run to bf16 rounding."""
    rec = load_golden(path)
    c = rec["case"]
    q = rec["q"].view(-1, c["hq"], c["d"])
    o = ao.decode_attention(q, rec["k_cache_after"], rec["v_cache_after"], rec["req_to_token"],
                            rec["req_pool_indices"], rec["seq_lens"], c["d"] ** -0.5, compute_dtype=torch.float32)
    assert (o.float().reshape(rec["o"].shape) - rec["o"].float()).abs().max() < 2e-2


def test_oracle_fp8_kv_write_matches_reference_golden():
    """oracle.set_kv_buffer_fp8 vs the reference's MHATokenToKVPool(dtype=float8_e4m3fn).set_kv_buffer, bit-exact,
    incl. the e4m3 edge values (tests/golden/make_golden_fp8kv.py)."""
    recs = load_golden(golden_files("kvwrite_fp8")[0])
    assert len(recs) >= 4
    for rec in recs:
        kc = torch.zeros_like(rec["k_cache_after"])
        vc = torch.zeros_like(rec["v_cache_after"])
        ao.set_kv_buffer_fp8(kc, vc, rec["loc"], rec["k"], rec["v"], rec["k_scale"], rec["v_scale"])
        assert torch.equal(kc, rec["k_cache_after"]) and torch.equal(vc, rec["v_cache_after"])
        # the edge rows really exercise the NaN code and the subnormal range
        assert (rec["k_cache_after"] == 0x7F).any() and (rec["k_cache_after"] == 0xFF).any()
