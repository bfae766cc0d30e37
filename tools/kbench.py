"""Development micro-benchmarks for single kernels (CUDA events, L2 flushed between iterations).

    python tools/kbench.py decode [--bs 64 --ctx 4096 --chunk 0]
    python tools/kbench.py gemm_int8 [--m 64]
Prints one JSON line per measurement.  bench.py is the judged entry point; this is the tuning loop.
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sglang_npu_b200 import kernels, synth  # noqa: E402
from sglang_npu_b200.srt.layers.attention.b200_backend import plan_decode_splits  # noqa: E402

PEAKS = {}
try:
    PEAKS = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
except Exception:
    pass
HBM = PEAKS.get("hbm_gbs", 6650.0)


def timeit(fn, iters=20, warmup=5, flush=True):
    dev = torch.device("cuda")
    scratch = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev) if flush else None
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    times = []
    for _ in range(iters):
        if flush:
            scratch.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        times.append(s.elapsed_time(e))
    times.sort()
    return times[len(times) // 2], times[0]


def bench_decode(args):
    dev = torch.device("cuda")
    bs, ctx, hq, hkv, d, page = args.bs, args.ctx, args.hq, args.hkv, args.d, args.page
    lens = [ctx] * bs if not args.ragged else torch.randint(ctx // 2, ctx + 1, (bs,), generator=torch.Generator().manual_seed(0)).tolist()
    pb = synth.make_paged_batch(lens, page, seed=1)
    k, v = synth.randn_kv_device(pb.num_slots, page, hkv, d, torch.bfloat16, dev)
    q = torch.randn(bs, hq, d, device=dev).to(torch.bfloat16)
    o = torch.empty_like(q)
    r2t = pb.req_to_token.to(dev)
    rpi = pb.req_pool_indices.to(dev)
    sl = pb.seq_lens.to(dev)
    kv_bytes = sum(lens) * 2 * hkv * d * 2
    for chunk in ([args.chunk] if args.chunk else [0, 256, 512, 1024, 2048]):
        ch, splits = plan_decode_splits(bs, hkv, hq // hkv, max(lens), 148, chunk)
        ws = torch.empty(max(kernels.decode_attn_workspace_bytes(bs, hq, d, splits), 16), dtype=torch.uint8, device=dev)
        fn = lambda: kernels.decode_attention_fwd(q, k, v, o, r2t, rpi, sl, ws, ch, splits, d ** -0.5)
        med, best = timeit(fn)
        print(json.dumps(dict(kernel="decode_attn", bs=bs, ctx=ctx, chunk=ch, splits=splits, ms_med=round(med, 4),
                              ms_best=round(best, 4), gbs_med=round(kv_bytes / med / 1e6, 1),
                              frac_of_measured=round(kv_bytes / med / 1e6 / HBM, 3))), flush=True)


def bench_gemm_int8(args):
    dev = torch.device("cuda")
    shapes = [(4096, 6144), (4096, 4096), (4096, 28672), (14336, 4096)]
    for m in ([args.m] if args.m else [1, 16, 64, 128, 256, 2048, 8192]):
        for (kk, n) in shapes:
            a = torch.randint(-127, 128, (m, kk), device=dev, dtype=torch.int8)
            w = torch.randint(-127, 128, (n, kk), device=dev, dtype=torch.int8)
            sa = torch.rand(m, device=dev)
            sb = torch.rand(n, device=dev)
            fn = lambda: kernels.int8_scaled_mm(a, w.t(), sa, sb, torch.bfloat16)
            med, best = timeit(fn)
            byts = n * kk + m * kk + m * n * 2
            print(json.dumps(dict(kernel="int8_scaled_mm", m=m, n=n, k=kk, ms_med=round(med, 4), ms_best=round(best, 4),
                                  gbs=round(byts / med / 1e6, 1), tops=round(2 * m * n * kk / med / 1e9, 1))), flush=True)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("what")
    ap.add_argument("--bs", type=int, default=64)
    ap.add_argument("--ctx", type=int, default=4096)
    ap.add_argument("--hq", type=int, default=32)
    ap.add_argument("--hkv", type=int, default=8)
    ap.add_argument("--d", type=int, default=128)
    ap.add_argument("--page", type=int, default=16)
    ap.add_argument("--chunk", type=int, default=0)
    ap.add_argument("--m", type=int, default=0)
    ap.add_argument("--ragged", action="store_true")
    args = ap.parse_args()
    {"decode": bench_decode, "gemm_int8": bench_gemm_int8}[args.what](args)
