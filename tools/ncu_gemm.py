"""Launch each Llama-3.1-8B decode GEMM shape twice at a given M (for `ncu -k regex:gemm_ -c 8`)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sglang_npu_b200 import kernels  # noqa: E402

m = int(sys.argv[1]) if len(sys.argv) > 1 else 64
cluster = len(sys.argv) > 2 and sys.argv[2] == "cluster"
dev = torch.device("cuda")
for rep in range(2):
    for (k, n) in [(4096, 6144), (4096, 4096), (4096, 28672), (14336, 4096)]:
        a = torch.randint(-127, 128, (m, k), device=dev, dtype=torch.int8)
        w = torch.randint(-127, 128, (n, k), device=dev, dtype=torch.int8)
        out = kernels.int8_scaled_mm(a, w.t(), torch.rand(m, device=dev), torch.rand(n, device=dev), torch.bfloat16,
                                     cluster_splitk=cluster)
        torch.cuda.synchronize()
print("ok")
