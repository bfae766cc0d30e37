"""Per-CTA trace rows of the block-FP8 stream-K GEMM (diagnostics): which CTAs finish last, and why.

    python tools/gemm_trace_cta.py [m] [n] [k]
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sglang_npu_b200 import kernels  # noqa: E402

GHZ = 1.965


def main():
    m, n, k = [int(x) for x in sys.argv[1:4]] if len(sys.argv) >= 4 else (64, 28672, 4096)
    dev = torch.device("cuda")
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    a8 = ((torch.rand(m, k, device=dev) - 0.5) * 2 * 448).to(torch.float8_e4m3fn)
    w8 = ((torch.rand(n, k, device=dev) - 0.5) * 2 * 448).to(torch.float8_e4m3fn)
    sa8 = torch.rand(m, k // 128, device=dev) * 1e-3
    sb8 = torch.rand(k // 128, n // 128, device=dev) * 1e-3
    fn = lambda: kernels.fp8_blockwise_scaled_mm(a8, w8.t(), sa8, sb8, torch.bfloat16)
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    buf = torch.zeros((1024, 16), dtype=torch.int64, device=dev)
    flush.zero_()
    torch.cuda.synchronize()
    kernels.set_trace(buf)
    fn()
    torch.cuda.synchronize()
    kernels.set_trace(None)
    t = buf.cpu()
    t = t[t[:, 0] != 0]
    G = t.shape[0]
    nkb = k // 128
    units = (n // 128) * nkb
    g0 = int(t[:, 0].min())
    rows = []
    for c in range(G):
        u0, u1 = c * units // G, (c + 1) * units // G
        rel = lambda i: int((t[c, i] - t[c, 14]).item() / GHZ) if t[c, i] != 0 else -1
        rows.append(dict(cta=c, kb0=u0 % nkb, units=u1 - u0, owner_wait=(u1 % nkb != 0), entry=int(t[c, 0]) - g0,
                         full0=rel(3), mma=rel(4), acc_last=rel(7), contrib=rel(6), flags=rel(8), epi=rel(9),
                         exit=int(t[c, 15]) - g0))
    rows.sort(key=lambda r: r["exit"])
    for r in rows[:5] + rows[len(rows) // 2 - 2:len(rows) // 2 + 3] + rows[-12:]:
        print(r)


if __name__ == "__main__":
    main()
