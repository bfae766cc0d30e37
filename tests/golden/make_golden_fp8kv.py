"""Golden vectors for the FP8 KV-cache write, produced by the UNMODIFIED reference on CPU (this container only).

    python tests/golden/make_golden_fp8kv.py

Runs the reference's own MHATokenToKVPool(dtype=float8_e4m3fn).set_kv_buffer (memory_pool.py:369-407, cast branch
:385-395) on seeded bf16/f16 rows — with and without k_scale / v_scale, including the e4m3 edge values (subnormals,
ties, +-448, the (448, 464] band that rounds down, > 464 -> NaN code, inf, NaN) — and stores inputs + the resulting
uint8 caches in kvwrite_fp8.pt.
"""
import os
import sys
import types

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import ref_shims  # noqa: E402

EDGE = [0.0, -0.0, 2.0 ** -10, 2.0 ** -9, 1.5 * 2.0 ** -9, 2.5 * 2.0 ** -9, 2.0 ** -7, 2.0 ** -6, 0.0146484375, 1.0625,
        1.1875, 17.0, 447.0, 448.0, 456.0, 464.0, 466.0, 472.0, 480.0, 1000.0, 65504.0, float("inf"), float("nan")]


def main():
    ref = ref_shims.load_reference_attention()
    Pool = ref["MHATokenToKVPool"]
    recs = []
    for dtype, hkv, d, T, ks, vs in [(torch.bfloat16, 2, 128, 37, None, None), (torch.bfloat16, 8, 128, 64, 0.37, 2.0),
                                     (torch.float16, 2, 64, 19, 1.7, None), (torch.float16, 1, 128, 5, None, 0.01)]:
        g = torch.Generator().manual_seed(1234)
        slots, page = 96, 16
        pool = Pool(size=slots, page_size=page, dtype=torch.float8_e4m3fn, head_num=hkv, head_dim=d, layer_num=1,
                    device="cpu", enable_memory_saver=False)
        k = (torch.randn((T, hkv, d), generator=g) * 3).to(dtype)
        v = (torch.randn((T, hkv, d), generator=g) * 40).to(dtype)
        edge = torch.tensor(EDGE + [-x for x in EDGE], dtype=torch.float32).to(dtype)
        k.view(-1)[: edge.numel()] = edge
        v.view(-1)[-edge.numel():] = edge
        loc = (torch.randperm(slots, generator=g)[:T] + page).long()
        layer = types.SimpleNamespace(layer_id=0)
        pool.set_kv_buffer(layer, loc, k.clone(), v.clone(), ks, vs)
        recs.append(dict(dtype=dtype, hkv=hkv, d=d, k=k, v=v, loc=loc, k_scale=ks, v_scale=vs, slots=slots, page=page,
                         k_cache_after=pool.k_buffer[0].clone(), v_cache_after=pool.v_buffer[0].clone()))
        assert pool.k_buffer[0].dtype == torch.uint8
    path = os.path.join(HERE, "kvwrite_fp8.pt")
    torch.save(recs, path)
    print(path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
