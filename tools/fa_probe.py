"""Bring-up probe for the two-tile TMA extend kernel: one small case against the oracle, errors per request printed."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import attention_oracle as ao  # noqa: E402
from sglang_npu_b200 import kernels, synth  # noqa: E402

lens = [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "300").split(",")]
prefix = [int(x) for x in (sys.argv[2] if len(sys.argv) > 2 else ",".join("0" for _ in lens)).split(",")]
page = int(sys.argv[3]) if len(sys.argv) > 3 else 16
hq, hkv, d = 8, 2, 128
dev = torch.device("cuda:0")
pb = synth.make_paged_batch(lens, page, seed=5)
k0, v0 = synth.randn_kv(pb.num_slots, page, hkv, d, torch.bfloat16, seed=15)
ext = [n - p for n, p in zip(lens, prefix)]
T = sum(ext)
q = torch.randn((T, hq, d), generator=torch.Generator().manual_seed(1)).to(torch.bfloat16)
o = torch.zeros_like(q, device=dev)
qo = torch.tensor([0] + list(torch.tensor(ext).cumsum(0)), dtype=torch.int32, device=dev)
pl = torch.tensor(prefix, dtype=torch.int32, device=dev)
kernels.extend_attention_fwd(q.to(dev), None, None, o, k0.to(dev), v0.to(dev), pb.req_to_token.to(dev),
                             pb.req_pool_indices.to(dev), qo, pl, max(ext), d ** -0.5, 0.0, True)
torch.cuda.synchronize()
ref = ao.extend_attention(q, k0, v0, pb.req_to_token, pb.req_pool_indices, pb.seq_lens, torch.tensor(prefix, dtype=torch.int32),
                          torch.tensor(ext, dtype=torch.int32), d ** -0.5, compute_dtype=torch.float32)
err = (o.cpu().float() - ref.float()).abs()
print("max err", err.max().item(), "nan", torch.isnan(o).sum().item())
start = 0
for i, e in enumerate(ext):
    blk = err[start:start + e]
    rows = blk.amax(dim=(1, 2))
    bad = (rows > 2e-2).nonzero().flatten()
    print(f"req {i}: ext {e} prefix {prefix[i]} max {blk.max().item():.4f} bad rows {bad.numel()} first {bad[:6].tolist()} "
          f"per-head max {[round(x, 3) for x in blk.amax(dim=(0, 2)).tolist()]}")
    start += e
