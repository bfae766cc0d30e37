"""Per-unit pipeline timeline of the fused AWQ GEMM (CTA 0): where a unit's ~N cycles go.  Diagnostic only.
    python tools/awq_trace.py [M] [K] [N]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sglang_npu_b200 import kernels, synth  # noqa: E402

m = int(sys.argv[1]) if len(sys.argv) > 1 else 128
kk = int(sys.argv[2]) if len(sys.argv) > 2 else 5120
n = int(sys.argv[3]) if len(sys.argv) > 3 else 55296
dev = torch.device("cuda")
qw, sc, qz = (t.to(dev) for t in synth.awq_weight(kk, n, 128, torch.float16, seed=1))
qwt, sz = kernels.awq_repack(qw, sc, qz)
x = torch.randn(m, kk, device=dev, dtype=torch.float16)
for _ in range(3):
    kernels.awq_gemm_repacked(x, qwt, sz, 128)
torch.cuda.synchronize()
units = 56
tb = torch.zeros((units, 16), dtype=torch.int64, device=dev)
kernels.set_trace(tb)
kernels.awq_gemm_repacked(x, qwt, sz, 128)
torch.cuda.synchronize()
kernels.set_trace(None)
t = tb.cpu()
names = ["dq0 w_full", "dq0 a_empty", "dq0 st issued", "dq0 a_full arrive", "dq1 w_full", "dq1 a_empty", "dq1 st issued",
         "dq1 a_full arrive", "mma x_full h0", "mma a_full c0", "mma a_full c1", "mma x_full h1", "mma issued", "tma W issue",
         "tma x0 issue", "tma x1 issue"]
t0 = t[:, 12].clone()  # MMA-issue-done of every unit as the per-unit clock
print("unit-to-unit period (MMA issue done), cycles:", (t0[8:48] - t0[7:47]).float().mean().item())
print("per unit, relative to the previous unit's 'mma issued' (mean over units 8..47):")
for k, nm in enumerate(names):
    rel = (t[8:48, k] - t0[7:47]).float()
    print(f"  {nm:20s} {rel.mean().item():9.0f}  (min {rel.min().item():7.0f} max {rel.max().item():7.0f})")
