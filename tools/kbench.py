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


def timeit_rot(make_fn, copies, iters=24):
    """Device time per launch with the launch queue kept full: a ~1 ms spin kernel runs while the host
    enqueues `iters` launches that rotate over `copies` independent operand sets (total bytes > L2, so
    every launch streams cold weights); CUDA events bracket the batch.  Host launch latency is excluded."""
    fns = [make_fn(c) for c in range(copies)]
    for f in fns:
        f()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda._sleep(3_000_000)
        s.record()
        for i in range(iters):
            fns[i % copies]()
        e.record()
        torch.cuda.synchronize()
        best = min(best, s.elapsed_time(e) / iters)
    return best


def bench_decode(args):
    dev = torch.device("cuda")
    bs, ctx, hq, hkv, d, page = args.bs, args.ctx, args.hq, args.hkv, args.d, args.page
    lens = [ctx] * bs if not args.ragged else torch.randint(ctx // 2, ctx + 1, (bs,), generator=torch.Generator().manual_seed(0)).tolist()
    pb = synth.make_paged_batch(lens, page, seed=1, shared_prefix=args.shared)
    k, v = synth.randn_kv_device(pb.num_slots, page, hkv, d, torch.bfloat16, dev)
    if args.fp8kv:
        k, v = k.to(torch.float8_e4m3fn), v.to(torch.float8_e4m3fn)
    q = torch.randn(bs, hq, d, device=dev).to(torch.bfloat16)
    o = torch.empty_like(q)
    r2t = pb.req_to_token.to(dev)
    rpi = pb.req_pool_indices.to(dev)
    sl = pb.seq_lens.to(dev)
    kv_bytes = sum(lens) * 2 * hkv * d * k.element_size()
    for chunk in ([args.chunk] if args.chunk else [0, 256, 512, 1024, 2048]):
        ch, splits = plan_decode_splits(bs, hkv, hq // hkv, max(lens), 148, chunk)
        splits += args.extra_splits  # a graph captured under a longer context bound: trailing splits exit at once
        ws = torch.empty(max(kernels.decode_attn_workspace_bytes(bs, hq, d, splits), 16), dtype=torch.uint8, device=dev)
        fn = lambda: kernels.decode_attention_fwd(q, k, v, o, r2t, rpi, sl, ws, ch, splits, d ** -0.5)
        if args.shared:  # shared-prefix (cascade) decode next to the plain kernel on the same batch
            uch, usp = plan_decode_splits(bs, hkv, hq // hkv, max(lens) - args.shared, 148, chunk)
            seg = kernels.plan_shared_prefix_segments(args.shared, bs, hq, head_dim=d)
            nseg = (args.shared + seg - 1) // seg
            ws2 = torch.empty(kernels.decode_attn_workspace_bytes(bs, hq, d, nseg + usp), dtype=torch.uint8, device=dev)
            if args.fp8kv:  # staged prefix (gather + up-cast, timed with the decode) for the tensor-core pass
                sk = torch.empty((args.shared, hkv, d), dtype=torch.bfloat16, device=dev)
                sv = torch.empty_like(sk)
                pl1 = torch.full((1,), args.shared, dtype=torch.int32, device=dev)
                off1 = torch.zeros(1, dtype=torch.int32, device=dev)
                tab = torch.arange(args.shared, dtype=torch.int32, device=dev)
                idx0 = torch.zeros(1, dtype=rpi.dtype, device=dev)

                def fn2():
                    kernels.fp8_kv_gather_dequant(sk, sv, k, v, r2t, rpi[:1], pl1, off1, args.shared)
                    kernels.decode_attention_shared_prefix_fwd(q, k, v, o, r2t, rpi, sl, args.shared, seg, ws2, uch, usp,
                                                               d ** -0.5, stage=(sk, sv, tab, idx0))
            else:
                fn2 = lambda: kernels.decode_attention_shared_prefix_fwd(q, k, v, o, r2t, rpi, sl, args.shared, seg, ws2,
                                                                         uch, usp, d ** -0.5)
            med2, best2 = timeit(fn2)
            unique = (args.shared + sum(n - args.shared for n in lens)) * 2 * hkv * d * k.element_size()
            print(json.dumps(dict(kernel="decode_attn_shared_prefix" + ("_fp8kv" if args.fp8kv else ""), bs=bs, ctx=ctx, shared=args.shared, seg_len=seg,
                                  segments=nseg, chunk=uch, splits=usp, ms_med=round(med2, 4), ms_best=round(best2, 4),
                                  unique_gbs=round(unique / med2 / 1e6, 1), logical_gbs=round(kv_bytes / med2 / 1e6, 1),
                                  prefix_tflops=round(4.0 * hq * d * bs * args.shared / med2 / 1e9, 1))), flush=True)
        med, best = timeit(fn)
        print(json.dumps(dict(kernel="decode_attn_fp8kv" if args.fp8kv else "decode_attn", bs=bs, ctx=ctx, chunk=ch, splits=splits, ms_med=round(med, 4),
                              ms_best=round(best, 4), gbs_med=round(kv_bytes / med / 1e6, 1),
                              frac_of_measured=round(kv_bytes / med / 1e6 / HBM, 3))), flush=True)


def bench_gemm_int8(args):
    dev = torch.device("cuda")
    tp = args.tp or 1  # Llama-3.1-8B (or --big: Llama-3.1-70B) shard shapes [K, N] of one TP rank
    shapes = [(4096, 6144 // tp), (4096 // tp, 4096), (4096, 28672 // tp), (14336 // tp, 4096)]
    if args.big:
        shapes = [(8192, 10240 // tp), (8192 // tp, 8192), (8192, 57344 // tp), (28672 // tp, 8192)]
    for m in ([args.m] if args.m else [1, 16, 64, 128, 256, 512, 2048, 8192]):
        tot = 0.0
        for (kk, n) in shapes:
            a = torch.randint(-127, 128, (m, kk), device=dev, dtype=torch.int8)
            copies = max(2, int(300e6 // (n * kk)) + 1)
            ws = [torch.randint(-127, 128, (n, kk), device=dev, dtype=torch.int8) for _ in range(copies)]
            sa = torch.rand(m, device=dev)
            sb = torch.rand(n, device=dev)
            med = best = timeit_rot(lambda c: (lambda: kernels.int8_scaled_mm(a, ws[c].t(), sa, sb, torch.bfloat16,
                                                                              cluster_splitk=args.cluster)), copies)
            byts = n * kk + m * kk + m * n * 2
            tot += med
            print(json.dumps(dict(kernel="int8_scaled_mm", path="cluster" if args.cluster else "streamk", tp=tp, sk_bn=os.environ.get("B200_SK_BN", "default"), m=m, n=n, k=kk, ms_med=round(med, 4), ms_best=round(best, 4),
                                  gbs=round(byts / med / 1e6, 1), tops=round(2 * m * n * kk / med / 1e9, 1))), flush=True)
        print(json.dumps(dict(kernel="int8_scaled_mm layer total", tp=tp, sk_bn=os.environ.get("B200_SK_BN", "default"), m=m, us=round(tot * 1e3, 1))), flush=True)


def bench_gemm_fp8(args):
    """Block-scaled (--block, default) or per-channel (--chan) FP8 GEMM at the Llama-8B shapes, next to cuBLASLt fp8
    (torch._scaled_mm, per-tensor scales: the library ceiling; blockwise scaling when this torch/cuBLASLt offers it)."""
    dev = torch.device("cuda")
    shapes = [(4096, 6144), (4096, 4096), (4096, 28672), (14336, 4096)]
    one = torch.ones((), device=dev)
    for m in ([args.m] if args.m else [64, 128, 2048, 8192]):
        for (kk, n) in shapes:
            a = ((torch.rand(m, kk, device=dev) - 0.5) * 2 * 448).to(torch.float8_e4m3fn)
            copies = max(2, int(300e6 // (n * kk)) + 1) if m <= 256 else 1
            ws = [((torch.rand(n, kk, device=dev) - 0.5) * 2 * 448).to(torch.float8_e4m3fn) for _ in range(copies)]
            byts = n * kk + m * kk + m * n * 2
            row = dict(kernel="fp8_blockwise_scaled_mm" if not args.chan else "fp8_scaled_mm", m=m, n=n, k=kk)
            if args.chan:
                sa, sb = torch.rand(m, device=dev) * 1e-3, torch.rand(n, device=dev) * 1e-3
                t = timeit_rot(lambda c: (lambda: kernels.fp8_scaled_mm(a, ws[c].t(), sa, sb, torch.bfloat16)), copies)
            else:
                sa = torch.rand(m, kk // 128, device=dev) * 1e-3
                sb = torch.rand(kk // 128, n // 128, device=dev) * 1e-3
                t = timeit_rot(lambda c: (lambda: kernels.fp8_blockwise_scaled_mm(a, ws[c].t(), sa, sb, torch.bfloat16)), copies)
            row.update(ms=round(t, 4), gbs=round(byts / t / 1e6, 1), frac_hbm=round(byts / t / 1e6 / HBM, 3),
                       tflops=round(2 * m * n * kk / t / 1e9, 1))
            try:
                t2 = timeit_rot(lambda c: (lambda: torch._scaled_mm(a, ws[c].t(), one, one, out_dtype=torch.bfloat16)), copies)
                row["cublaslt_fp8_tensorwise_ms"] = round(t2, 4)
                row["cublaslt_fp8_tensorwise_tflops"] = round(2 * m * n * kk / t2 / 1e9, 1)
                row["vs_cublaslt"] = round(t2 / t, 3)
            except Exception as e:  # noqa: BLE001
                row["cublaslt_err"] = str(e)[:80]
            if not args.chan:
                try:  # blockwise 1x128 x 128x128 scaling inside cuBLASLt, where supported
                    sa_c = sa.t().contiguous().t()
                    sb_c = sb.t().contiguous()
                    t3 = timeit_rot(lambda c: (lambda: torch._scaled_mm(a, ws[c].t(), sa_c, sb_c.t(), out_dtype=torch.bfloat16)), copies)
                    row["cublaslt_fp8_blockwise_ms"] = round(t3, 4)
                except Exception as e:  # noqa: BLE001
                    row["cublaslt_blockwise_err"] = str(e)[:60]
            print(json.dumps(row), flush=True)


def bench_awq(args):
    """cfg 4 (Qwen2.5-32B) per-GPU shapes: decode M = 128, prefill M = 2048."""
    dev = torch.device("cuda")
    hidden, inter, hq, hkv, d = 5120, 27648, 40, 8, 128
    for tp in ([args.tp] if args.tp else [1, 2, 8]):
        n_qkv = (hq // tp + 2 * max(1, hkv // tp)) * d
        shapes = [(hidden, n_qkv), (hq * d // tp, hidden), (hidden, 2 * inter // tp), (inter // tp, hidden)]
        for m in ([args.m] if args.m else [128, 2048]):
            tot_t, tot_b = 0.0, 0
            for (kk, n) in shapes:
                qw, sc, qz = synth.awq_weight(kk, n, 128, torch.float16, seed=1)
                qw, sc, qz = qw.to(dev), sc.to(dev), qz.to(dev)
                copies = max(2, int(300e6 // (qw.numel() * 4)) + 1)
                qwt, sz = kernels.awq_repack(qw, sc, qz)  # streaming layout, produced once at weight-loading time
                qws = [qwt.clone() for _ in range(copies)]
                x = torch.randn(m, kk, device=dev, dtype=torch.float16)
                med = best = timeit_rot(lambda c: (lambda: kernels.awq_gemm_repacked(x, qws[c], sz, 128)), copies)
                byts = qw.numel() * 4 + sc.numel() * 2 + qz.numel() * 4
                tot_t += med
                tot_b += byts
                print(json.dumps(dict(kernel="awq_gemm", tp=tp, m=m, n=n, k=kk, ms_med=round(med, 4), ms_best=round(best, 4),
                                      gbs=round(byts / med / 1e6, 1), frac_hbm=round(byts / med / 1e6 / HBM, 3),
                                      tflops=round(2 * m * n * kk / med / 1e9, 1))), flush=True)
            print(json.dumps(dict(kernel="awq_gemm layer total", tp=tp, m=m, ms=round(tot_t, 4),
                                  gbs=round(tot_b / tot_t / 1e6, 1), frac_hbm=round(tot_b / tot_t / 1e6 / HBM, 3))), flush=True)


def bench_extend(args):
    """cfg 3 layer shape: chunked prefill of an 8k prompt (chunks of `--chunk` tokens, prefix = what is cached)."""
    dev = torch.device("cuda")
    hq, hkv, d, page = args.hq, args.hkv, args.d, args.page
    total = args.ctx
    chunk = args.chunk or total
    pb = synth.make_paged_batch([total], page, seed=1)
    kc, vc = synth.randn_kv_device(pb.num_slots, page, hkv, d, torch.bfloat16, dev)
    r2t, rpi = pb.req_to_token.to(dev), pb.req_pool_indices.to(dev)
    for prefix in range(0, total, chunk):
        e = min(chunk, total - prefix)
        q = torch.randn(e, hq, d, device=dev).to(torch.bfloat16)
        o = torch.empty_like(q)
        qo = torch.tensor([0, e], dtype=torch.int32, device=dev)
        pl = torch.tensor([prefix], dtype=torch.int32, device=dev)
        fn = lambda: kernels.extend_attention_fwd(q, None, None, o, kc, vc, r2t, rpi, qo, pl, e, d ** -0.5, 0.0, True)
        med, best = timeit(fn, iters=10, warmup=3, flush=False)
        flops = 4 * d * hq * (e * prefix + e * (e + 1) / 2)
        print(json.dumps(dict(kernel="extend_attn", legacy=os.environ.get("B200_EXTEND_LEGACY", "0"), extend=e, prefix=prefix,
                              ms_med=round(med, 4), tflops=round(flops / med / 1e9, 1))), flush=True)


def bench_peak(args):
    """Library ceilings measured the way MEASURED_PEAKS.json measures bf16: cuBLASLt int8 (torch._int_mm) and
    fp8 (torch._scaled_mm) at 8192^3 and at the Llama-8B prefill shapes, plus SM clocks under load."""
    import subprocess
    dev = torch.device("cuda")
    shapes = [(8192, 8192, 8192), (8192, 6144, 4096), (8192, 28672, 4096), (8192, 4096, 14336)]
    for (m, n, k) in shapes:
        a = torch.randint(-127, 128, (m, k), device=dev, dtype=torch.int8)
        b = torch.randint(-127, 128, (n, k), device=dev, dtype=torch.int8)
        row = dict(m=m, n=n, k=k)
        try:
            t = timeit_rot(lambda c: (lambda: torch._int_mm(a, b.t())), 1, iters=16)
            row["cublaslt_int8_tops"] = round(2 * m * n * k / t / 1e9, 1)
        except Exception as e:  # noqa: BLE001
            row["cublaslt_int8_err"] = str(e)[:80]
        try:
            a8, b8 = a.to(torch.float8_e4m3fn), b.to(torch.float8_e4m3fn)
            one = torch.ones((), device=dev)
            t = timeit_rot(lambda c: (lambda: torch._scaled_mm(a8, b8.t(), one, one, out_dtype=torch.bfloat16)), 1, iters=16)
            row["cublaslt_fp8_tflops"] = round(2 * m * n * k / t / 1e9, 1)
        except Exception as e:  # noqa: BLE001
            row["cublaslt_fp8_err"] = str(e)[:80]
        sa, sb = torch.rand(m, device=dev), torch.rand(n, device=dev)
        t = timeit_rot(lambda c: (lambda: kernels.int8_scaled_mm(a, b.t(), sa, sb, torch.bfloat16)), 1, iters=16)
        row["b200_int8_tops"] = round(2 * m * n * k / t / 1e9, 1)
        q = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader"], capture_output=True, text=True)
        row["smi_after"] = q.stdout.strip()
        print(json.dumps(row), flush=True)


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
    ap.add_argument("--tp", type=int, default=0, help="awq: one TP degree instead of 1, 2 and 8")
    ap.add_argument("--extra-splits", type=int, default=0, dest="extra_splits",
                    help="decode: over-provision the split grid (a graph captured under a longer context bound)")
    ap.add_argument("--fp8kv", action="store_true", help="decode: FP8 (e4m3) KV cache")
    ap.add_argument("--shared", type=int, default=0, help="decode: every request shares this many leading tokens")
    ap.add_argument("--big", action="store_true", help="gemm_int8: Llama-3.1-70B shapes instead of 8B")
    ap.add_argument("--cluster", action="store_true", help="gemm_int8: cluster split-K kernel instead of stream-K")
    ap.add_argument("--chan", action="store_true", help="gemm_fp8: per-channel scales (fp8_scaled_mm) instead of block scales")
    args = ap.parse_args()
    {"decode": bench_decode, "gemm_int8": bench_gemm_int8, "awq": bench_awq, "peak": bench_peak, "extend": bench_extend,
     "gemm_fp8": bench_gemm_fp8}[args.what](args)
