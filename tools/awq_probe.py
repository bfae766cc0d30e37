"""One AWQ GEMM shape launched a few times — the command profiled under `ncu --set full`.
    python tools/awq_probe.py [M] [K] [N]"""
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
print("ok")
