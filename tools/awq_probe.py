"""One AWQ GEMM shape launched a few times — the command profiled under `ncu --set full`."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sglang_npu_b200 import kernels, synth
m = int(sys.argv[1]) if len(sys.argv) > 1 else 128
kk, n = 5120, 55296
dev = torch.device("cuda")
qw, sc, qz = (t.to(dev) for t in synth.awq_weight(kk, n, 128, torch.float16, seed=1))
x = torch.randn(m, kk, device=dev, dtype=torch.float16)
for _ in range(3):
    kernels.awq_gemm(x, qw, sc, qz)
torch.cuda.synchronize()
print("ok")
