"""One block-FP8 prefill GEMM per shape at M = 8192 (for `ncu -k regex:gemm_block_prefill`)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sglang_npu_b200 import kernels  # noqa: E402

m = 8192
dev = torch.device("cuda")
for (k, n) in [(4096, 28672), (14336, 4096)]:
    a = ((torch.rand(m, k, device=dev) - 0.5) * 2 * 448).to(torch.float8_e4m3fn)
    w = ((torch.rand(n, k, device=dev) - 0.5) * 2 * 448).to(torch.float8_e4m3fn)
    sa = torch.rand(m, k // 128, device=dev) * 1e-3
    sb = torch.rand(k // 128, n // 128, device=dev) * 1e-3
    for _ in range(2):
        kernels.fp8_blockwise_scaled_mm(a, w.t(), sa, sb, torch.bfloat16)
    torch.cuda.synchronize()
print("ok")
