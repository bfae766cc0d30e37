"""One int8 prefill GEMM launch per Llama-3.1-8B shape at M = 8192 (for `ncu -k regex:gemm_prefill -c 4`)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sglang_npu_b200 import kernels  # noqa: E402

m = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
dev = torch.device("cuda")
for (k, n) in [(4096, 28672), (14336, 4096)]:
    a = torch.randint(-127, 128, (m, k), device=dev, dtype=torch.int8)
    w = torch.randint(-127, 128, (n, k), device=dev, dtype=torch.int8)
    for _ in range(2):
        kernels.int8_scaled_mm(a, w.t(), torch.rand(m, device=dev), torch.rand(n, device=dev), torch.bfloat16)
    torch.cuda.synchronize()
print("ok")
