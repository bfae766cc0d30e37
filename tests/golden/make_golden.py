"""Generate golden vectors by running the UNMODIFIED reference on CPU (this container only).

    python tests/golden/make_golden.py

Imports TorchNativeAttnBackend / RadixAttention / MHATokenToKVPool / ReqToTokenPool / ForwardBatch from
/root/reference/python through oracle/ref_shims.py, drives them exactly like the reference's own
backend unit test does (python/sglang/test/attention/test_flashattn_backend.py:15-69, :163-200:
MockModelRunner + hand-built ForwardBatch) on seeded synthetic inputs, and stores inputs + outputs
as small .pt fixtures next to this script.  The fixtures travel to the GPU box; the reference does not.
"""
import os
import sys
import types

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import ref_shims  # noqa: E402
from sglang_npu_b200 import synth  # noqa: E402

CASES = [
    # name, Hq, Hkv, D, page, seq_lens, prefix_lens (extend case), dtype
    dict(name="decode_gqa4_d128_p16", mode="decode", hq=8, hkv=2, d=128, page=16, lens=[37, 16, 1, 130], dtype="bf16"),
    dict(name="decode_gqa7_d64_p1", mode="decode", hq=14, hkv=2, d=64, page=1, lens=[5, 64, 33], dtype="bf16"),
    dict(name="decode_mha_d128_p16_f16", mode="decode", hq=4, hkv=4, d=128, page=16, lens=[48, 17], dtype="f16"),
    dict(name="extend_prefix_d128_p16", mode="extend", hq=8, hkv=2, d=128, page=16, lens=[70, 33, 16],
         prefix=[32, 0, 15], dtype="bf16"),
    dict(name="extend_noprefix_d64_p1", mode="extend", hq=14, hkv=2, d=64, page=1, lens=[9, 65], prefix=[0, 0],
         dtype="bf16"),
]

DT = {"bf16": torch.bfloat16, "f16": torch.float16}


def run_case(ref, c):
    dtype = DT[c["dtype"]]
    torch.manual_seed(1234)
    pb = synth.make_paged_batch(c["lens"], c["page"], seed=1234)
    bs = len(c["lens"])
    r2t_pool = ref["ReqToTokenPool"](size=pb.req_to_token.shape[0], max_context_len=pb.req_to_token.shape[1],
                                     device="cpu", enable_memory_saver=False)
    r2t_pool.req_to_token.copy_(pb.req_to_token)
    kv_pool = ref["MHATokenToKVPool"](size=pb.num_slots, page_size=c["page"], dtype=dtype, head_num=c["hkv"],
                                      head_dim=c["d"], layer_num=1, device="cpu", enable_memory_saver=False)
    k0, v0 = synth.randn_kv(pb.num_slots, c["page"], c["hkv"], c["d"], dtype, seed=7)
    kv_pool.k_buffer[0].copy_(k0)
    kv_pool.v_buffer[0].copy_(v0)
    layer = ref["RadixAttention"](num_heads=c["hq"], head_dim=c["d"], scaling=c["d"] ** -0.5,
                                  num_kv_heads=c["hkv"], layer_id=0)
    runner = types.SimpleNamespace(device="cpu")
    backend = ref["TorchNativeAttnBackend"](runner)
    FM = ref["ForwardMode"]
    seq_lens = pb.seq_lens
    if c["mode"] == "decode":
        T = bs
        out_loc = torch.stack([pb.req_to_token[pb.req_pool_indices[i], seq_lens[i] - 1] for i in range(bs)]).long()
        fb = ref["ForwardBatch"](forward_mode=FM.DECODE, batch_size=bs, input_ids=torch.zeros(T, dtype=torch.long),
                                 req_pool_indices=pb.req_pool_indices, seq_lens=seq_lens, out_cache_loc=out_loc,
                                 seq_lens_sum=int(seq_lens.sum()), attn_backend=backend)
        extra = {}
    else:
        prefix = torch.tensor(c["prefix"], dtype=torch.int32)
        ext = (seq_lens - prefix).to(torch.int32)
        T = int(ext.sum())
        out_loc = torch.cat([pb.req_to_token[pb.req_pool_indices[i], int(prefix[i]) : int(seq_lens[i])]
                             for i in range(bs)]).long()
        fb = ref["ForwardBatch"](forward_mode=FM.EXTEND, batch_size=bs, input_ids=torch.zeros(T, dtype=torch.long),
                                 req_pool_indices=pb.req_pool_indices, seq_lens=seq_lens, out_cache_loc=out_loc,
                                 seq_lens_sum=int(seq_lens.sum()), attn_backend=backend,
                                 extend_num_tokens=T, extend_seq_lens=ext, extend_prefix_lens=prefix,
                                 extend_prefix_lens_cpu=prefix.tolist(), extend_seq_lens_cpu=ext.tolist())
        extra = dict(extend_prefix_lens=prefix, extend_seq_lens=ext)
    fb.req_to_token_pool = r2t_pool
    fb.token_to_kv_pool = kv_pool
    g = torch.Generator().manual_seed(99)
    q = torch.randn((T, c["hq"] * c["d"]), generator=g).to(dtype)
    k = torch.randn((T, c["hkv"] * c["d"]), generator=g).to(dtype)
    v = torch.randn((T, c["hkv"] * c["d"]), generator=g).to(dtype)
    backend.init_forward_metadata(fb)
    o = layer(q.clone(), k.clone(), v.clone(), fb)
    return dict(case=c, req_to_token=pb.req_to_token, req_pool_indices=pb.req_pool_indices, seq_lens=seq_lens,
                out_cache_loc=out_loc, num_slots=pb.num_slots, k_cache_before=k0, v_cache_before=v0,
                q=q, k=k, v=v, o=o, k_cache_after=kv_pool.k_buffer[0].clone(),
                v_cache_after=kv_pool.v_buffer[0].clone(), **extra)


def main():
    ref = ref_shims.load_reference_attention()
    for c in CASES:
        rec = run_case(ref, c)
        path = os.path.join(HERE, f"attn_{c['name']}.pt")
        torch.save(rec, path)
        print(f"{path}: {os.path.getsize(path) / 1024:.1f} KiB, o abs-mean {rec['o'].float().abs().mean():.4f}")


if __name__ == "__main__":
    main()
