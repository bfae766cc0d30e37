"""Block-FP8 decode GEMM (stream-K weight streaming) on the gate_up shape at M = 64, a few launches
(for `ncu -k regex:gemm_block_streamk`)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sglang_npu_b200 import kernels  # noqa: E402

m = int(sys.argv[1]) if len(sys.argv) > 1 else 64
dev = torch.device("cuda")
k, n = 4096, 28672
a = ((torch.rand(m, k, device=dev) - 0.5) * 2 * 448).to(torch.float8_e4m3fn)
w = ((torch.rand(n, k, device=dev) - 0.5) * 2 * 448).to(torch.float8_e4m3fn)
sa = torch.rand(m, k // 128, device=dev) * 1e-3
sb = torch.rand(k // 128, n // 128, device=dev) * 1e-3
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
for _ in range(4):
    flush.zero_()
    kernels.fp8_blockwise_scaled_mm(a, w.t(), sa, sb, torch.bfloat16)
    torch.cuda.synchronize()
print("ok")
