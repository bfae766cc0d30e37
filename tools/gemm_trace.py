"""Per-CTA phase timeline of the decode GEMM kernels (b200_debug_set_trace).  Diagnostics only.

    python tools/gemm_trace.py [m ...]
For each Llama-3.1-8B GEMM shape and each kernel (stream-K, cluster split-K) prints the kernel span seen by the
CTAs (globaltimer) and median / max over CTAs of the phase boundaries, in ns relative to CTA entry.
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sglang_npu_b200 import kernels  # noqa: E402

GHZ = 1.965
NAMES = {
    "streamk": {1: "setup", 2: "pdl_wait", 3: "first_full", 4: "mma_issued", 5: "seg0_acc", 6: "contrib_pub",
                7: "last_seg_acc", 8: "flags_seen", 11: "partials_in_smem", 9: "epi_done"},
    "block": {1: "setup", 2: "pdl_wait", 12: "scales_ready", 3: "first_full", 4: "mma_issued", 5: "unit0_acc",
              6: "contrib_pub", 7: "last_unit_acc", 8: "flags_seen", 9: "epi_done"},
    "cluster": {1: "setup", 2: "pdl_wait", 3: "first_full", 4: "mma_issued", 5: "acc_ready", 6: "csync1", 7: "scatter",
                8: "csync2", 9: "epi_done"},
}


def main():
    ms = [int(x) for x in sys.argv[1:] if x.isdigit()] or [16, 64]
    fp8_block = "block" in sys.argv
    dev = torch.device("cuda")
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    for m in ms:
        for (k, n) in [(4096, 6144), (4096, 4096), (4096, 28672), (14336, 4096)]:
            a = torch.randint(-127, 128, (m, k), device=dev, dtype=torch.int8)
            w = torch.randint(-127, 128, (n, k), device=dev, dtype=torch.int8)
            sa, sb = torch.rand(m, device=dev), torch.rand(n, device=dev)
            if fp8_block:
                a8 = ((torch.rand(m, k, device=dev) - 0.5) * 2 * 448).to(torch.float8_e4m3fn)
                w8 = ((torch.rand(n, k, device=dev) - 0.5) * 2 * 448).to(torch.float8_e4m3fn)
                sa8 = torch.rand(m, k // 128, device=dev) * 1e-3
                sb8 = torch.rand(k // 128, n // 128, device=dev) * 1e-3
            for path in (("streamk", "block") if fp8_block else ("streamk", "cluster")):
                fn = lambda: kernels.int8_scaled_mm(a, w.t(), sa, sb, torch.bfloat16, cluster_splitk=(path == "cluster"))
                if path == "block":
                    fn = lambda: kernels.fp8_blockwise_scaled_mm(a8, w8.t(), sa8, sb8, torch.bfloat16)
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
                g0 = int(t[:, 0].min())
                row = dict(m=m, n=n, k=k, path=path, ctas=int(t.shape[0]),
                           span_ns=int(t[:, 15].max()) - g0,
                           entry_spread_ns=int(t[:, 0].max()) - g0,
                           exit_med_ns=int((t[:, 15] - g0).median()))
                for idx, name in NAMES[path].items():
                    col = t[:, idx]
                    ok = col != 0
                    if ok.any():
                        d = (col[ok] - t[ok, 14]).float() / GHZ
                        row[name] = [int(d.median()), int(d.max())]
                if path == "streamk" and not fp8_block:
                    row["peers_max"] = int(t[:, 10].max())
                print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
