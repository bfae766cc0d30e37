"""Does programmatic dependent launch (PDL) survive torch's stream capture, and what does it buy?
    B200_PDL=1 python tools/pdl_probe.py ; B200_PDL=0 python tools/pdl_probe.py
Times chains of [per_token_quant_int8 -> int8_scaled_mm] eagerly and as a CUDA graph, dumps the graph as DOT."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sglang_npu_b200 import kernels  # noqa: E402

dev = torch.device("cuda")
M, K, N = 64, 4096, 6144
L = 32
x = torch.randn(M, K, device=dev).bfloat16()
ws = [torch.randint(-127, 128, (N, K), device=dev, dtype=torch.int8) for _ in range(L)]
sb = torch.rand(N, device=dev)


def chain_qg():
    for w in ws:
        xq, xs = kernels.per_token_quant_int8(x)
        kernels.int8_scaled_mm(xq, w.t(), xs, sb, torch.bfloat16)


xq0, xs0 = kernels.per_token_quant_int8(x)


def chain_g():
    for w in ws:
        kernels.int8_scaled_mm(xq0, w.t(), xs0, sb, torch.bfloat16)


def time_fn(fn, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(iters):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda._sleep(2_000_000)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        best = min(best, s.elapsed_time(e))
    return best


out = {"pdl": os.environ.get("B200_PDL", "1")}
for name, fn in (("quant+gemm", chain_qg), ("gemm", chain_g)):
    out[f"eager_{name}_us_per_iter"] = round(time_fn(fn) * 1e3 / L, 2)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        fn()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    out[f"graph_{name}_us_per_iter"] = round(time_fn(g.replay) * 1e3 / L, 2)
print(json.dumps(out))
