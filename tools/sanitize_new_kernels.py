"""Small launches of the round-2 kernels for compute-sanitizer (memcheck / racecheck / synccheck are too slow for the
full suite): two-tile TMA extend attention (page 16 and page 1, ragged), pair GEMM (odd tile counts), token-block
stream-K, FP8-KV shared-prefix decode."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sglang_npu_b200 import kernels, synth  # noqa: E402

dev = torch.device("cuda:0")
hq, hkv, d = 8, 2, 128
for page in (16, 1):
    lens, prefix = [300, 513], [16, 0]
    pb = synth.make_paged_batch(lens, page, seed=5)
    k0, v0 = synth.randn_kv(pb.num_slots, page, hkv, d, torch.bfloat16, seed=15)
    ext = [n - p for n, p in zip(lens, prefix)]
    q = torch.randn((sum(ext), hq, d)).to(torch.bfloat16).to(dev)
    o = torch.zeros_like(q)
    qo = torch.tensor([0, ext[0], ext[0] + ext[1]], dtype=torch.int32, device=dev)
    kernels.extend_attention_fwd(q, None, None, o, k0.to(dev), v0.to(dev), pb.req_to_token.to(dev), pb.req_pool_indices.to(dev),
                                 qo, torch.tensor(prefix, dtype=torch.int32, device=dev), max(ext), d ** -0.5, 0.0, True)
    torch.cuda.synchronize()
for (m, n, k) in [(257, 512, 256), (640, 264, 1040), (300, 768, 512)]:
    a = torch.randint(-127, 128, (m, k), device=dev, dtype=torch.int8)
    w = torch.randint(-127, 128, (n, k), device=dev, dtype=torch.int8)
    kernels.int8_scaled_mm(a, w.t(), torch.rand(m, device=dev), torch.rand(n, device=dev), torch.bfloat16)
    torch.cuda.synchronize()
# FP8-KV shared-prefix decode
lens = [600, 700, 513]
pb = synth.make_paged_batch(lens, 16, seed=7, shared_prefix=512)
k, v = synth.randn_kv_device(pb.num_slots, 16, hkv, d, torch.bfloat16, dev, seed=8)
k8, v8 = k.to(torch.float8_e4m3fn).view(torch.uint8), v.to(torch.float8_e4m3fn).view(torch.uint8)
bs = len(lens)
q = torch.randn(bs, hq, d, device=dev).to(torch.bfloat16)
o = torch.empty_like(q)
r2t, rpi, sl = pb.req_to_token.to(dev), pb.req_pool_indices.to(dev), pb.seq_lens.to(dev)
seg = kernels.plan_shared_prefix_segments(512, bs, hq)
nseg = (512 + seg - 1) // seg
ws = torch.empty(kernels.decode_attn_workspace_bytes(bs, hq, d, nseg + 3), dtype=torch.uint8, device=dev)
sk = torch.empty((512, hkv, d), dtype=torch.bfloat16, device=dev)
sv = torch.empty_like(sk)
kernels.fp8_kv_gather_dequant(sk, sv, k8, v8, r2t, rpi[:1], torch.full((1,), 512, dtype=torch.int32, device=dev),
                              torch.zeros(1, dtype=torch.int32, device=dev), 512)
kernels.decode_attention_shared_prefix_fwd(q, k8, v8, o, r2t, rpi, sl, 512, seg, ws, 256, 3, d ** -0.5,
                                           stage=(sk, sv, torch.arange(512, dtype=torch.int32, device=dev),
                                                  torch.zeros(1, dtype=rpi.dtype, device=dev)))
torch.cuda.synchronize()
print("ok")
