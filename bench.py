#!/usr/bin/env python
"""bench.py — the hot path's headline benchmark and its variants.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]           # the BASELINE line
    python bench.py --workload prefill [--quant fp8|int8] [...]                       # config 3 (non-headline)
    python bench.py --model qwen2.5-32b --quant awq --bs 128 --ctx 8192 --fixed-global-batch   (under torchrun, cfg 4)
    python bench.py --model llama-3.1-70b --bs 256 --ctx 32768 --shared-prefix 30720 --fixed-global-batch  (cfg 5)

DEFAULT (no workload flags) = BASELINE.json configs[1]: decode tok/s for Llama-3.1-8B W8A8-int8, bs 64 per GPU, 4k
context, page_size 16, bf16 KV.  Any other flag combination prints the same JSON shape with a `variant` key.

decode workload
* a "step" = one decode forward of the whole layer stack (embedding + N layers + lm_head + greedy argmax) over one
  synthetic batch; CUDA-graph replay, inputs resident in HBM -> `value` (whole-job tok/s).
* `e2e` = a real decode loop through the public call HotPathRunner.decode_loop_step() with PINNED HOST tensors: per
  step H2D of input_ids / req_pool_indices / seq_lens, the KV slot of the new token allocated on the device by the
  fused get_last_loc + alloc_decode + write_req_to_token_pool launch, graph replay, D2H of the sampled ids, which are
  the next step's input_ids; seq_lens advance by one per step (ctx+1, ctx+2, ...).
* N > 1: tensor parallel over N GPUs of one box.  Default weak scaling: global batch = bs * N so per-GPU KV bytes stay
  fixed; `--fixed-global-batch` keeps the global batch (configs 4 / 5 are quoted that way).

prefill workload (`--workload prefill`, BASELINE configs[2])
* a "step" = one chunked-prefill forward: `--reqs` requests x `--new` new tokens each (default 4 x 2048 = an
  8192-token chunk) over a `--prefix`-token radix-cache hit per request (default 6144), eager launches; value =
  new tokens / s.  roofline = the largest GEMM (gate_up) against a tensor-pipe peak measured in the same run.

`--impl reference`: the reference's own CPU implementation of the same path (oracle/_ref, compiled from
/root/reference/sgl-kernel/csrc/cpu) on all host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

UNIT = "tok/s"
PAGE = 16
DEFAULTS = dict(workload="decode", model="llama-3.1-8b", quant="int8", bs=64, ctx=4096, kv_dtype="bf16", shared_prefix=0,
                fixed_global_batch=False)
HEADLINE_METRIC = "decode tok/s (Llama-3.1-8B W8A8-int8, bs64/GPU, 4k ctx, page16)"
QUANT = {  # --quant -> (registry name, activation dtype, label, arithmetic dtype of the linears)
    "int8": ("b200_w8a8_int8", torch.bfloat16, "W8A8-int8", "int8"),
    "fp8": ("b200_fp8", torch.bfloat16, "W8A8-FP8 block-scaled [128,128]", "fp8_e4m3"),
    "fp8_chan": ("b200_w8a8_fp8", torch.bfloat16, "W8A8-FP8 per-token x per-channel", "fp8_e4m3"),
    "awq": ("b200_awq", torch.float16, "AWQ-int4 (group 128)", "int4->f16"),
}
QCFG = {"weight_block_size": [128, 128], "activation_scheme": "dynamic", "quant_method": "fp8", "bits": 4,
        "group_size": 128, "zero_point": True, "w_bit": 4, "q_group_size": 128}
MODEL_LABEL = {"llama-3.1-8b": "Llama-3.1-8B", "qwen2.5-32b": "Qwen2.5-32B", "llama-3.1-70b": "Llama-3.1-70B",
               "qwen2.5-0.5b": "Qwen2.5-0.5B"}


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return json.load(fh), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def metric_name(args) -> str:
    if is_headline(args):
        return HEADLINE_METRIC
    label = f"{MODEL_LABEL.get(args.model, args.model)} {QUANT[args.quant][2]}"
    if args.workload == "prefill":
        return (f"prefill tok/s ({label}, {args.reqs} x {args.new}-token chunk, {args.prefix}-token radix prefix, "
                f"page{PAGE})")
    per = "" if args.fixed_global_batch else "/GPU"
    sp = f", {args.shared_prefix}-token shared prefix" if args.shared_prefix else ""
    kv = ", fp8_e4m3 KV" if args.kv_dtype == "fp8" else ""
    return f"decode tok/s ({label}, bs{args.bs}{per}, {args.ctx} ctx{sp}{kv}, page{PAGE})"


def is_headline(args) -> bool:
    return all(getattr(args, k) == v for k, v in DEFAULTS.items())


def linear_shapes(cfg, tp: int = 1):
    """Per-layer GEMMs [K -> N] on one TP rank: qkv, o, gate_up, down."""
    hq, hkv, d = cfg.num_attention_heads // tp, cfg.get_num_kv_heads(tp), cfg.head_dim
    return [(cfg.hidden_size, (hq + 2 * hkv) * d), (hq * d, cfg.hidden_size),
            (cfg.hidden_size, 2 * cfg.intermediate_size // tp), (cfg.intermediate_size // tp, cfg.hidden_size)]


# ---------------------------------------------------------------------------------------------------------
# reference arm / CPU baseline: the reference's own CPU kernels on layers of the same workload
# ---------------------------------------------------------------------------------------------------------
def _cpu_kv(slots, hkv, d):
    blk = torch.randn(1 << 22, generator=torch.Generator().manual_seed(0)).to(torch.bfloat16)
    n = slots * hkv * d
    k = blk.repeat((n + blk.numel() - 1) // blk.numel())[:n].view(slots, hkv, d).contiguous()
    return k, k.flip(0).contiguous()


def _cpu_linears(ops, cfg, quant: str, M: int, dtype):
    """Returns a closure running the four linears of one layer with the reference's CPU kernels (or the oracle port)."""
    shapes = linear_shapes(cfg)
    xs = [(torch.randn(M, kk) * 0.1).to(dtype) for (kk, _) in shapes]
    if quant in ("fp8", "fp8_chan"):
        ws = []
        for i, (kk, nn_) in enumerate(shapes):
            g = torch.Generator().manual_seed(i)
            w = ((torch.rand((nn_, kk), generator=g) - 0.5) * 2 * 448.0).to(torch.float8_e4m3fn)
            s = torch.rand((nn_ // 128, kk // 128), generator=g) * 1e-3
            ws.append((w, s))
        if ops is not None:
            packed = [(ops.convert_weight_packed(w), s) for (w, s) in ws]
            # the reference's CPU block-FP8 linear is W8A16 (gemm_fp8.cpp:481-505): bf16 activations, e4m3 weights
            return (lambda: [ops.fp8_scaled_mm_cpu(x, pw, s, [128, 128], None, dtype, True)
                             for (pw, s), x in zip(packed, xs)]), "fp8_scaled_mm_cpu (W8A16)"
        from oracle import quant_oracle as qo

        return (lambda: [qo.w8a8_block_fp8_linear(x, w, s) for (w, s), x in zip(ws, xs)]), "oracle port block-fp8"
    if quant == "awq":
        # the reference has no CPU AWQ kernel: its CPU path dequantises and calls the bf16/fp16 linear; timed as the
        # oracle port (awq_dequantize_torch + matmul)
        from oracle import quant_oracle as qo
        from sglang_npu_b200 import synth

        ws = [synth.awq_weight(kk, nn_, 128, torch.float16, seed=i) for i, (kk, nn_) in enumerate(shapes)]
        return (lambda: [qo.awq_linear(x.half(), qw, sc, qz, 128) for (qw, sc, qz), x in zip(ws, xs)]), \
            "oracle port awq dequant + matmul"
    ws = []
    for i, (kk, nn_) in enumerate(shapes):
        w = torch.randint(-127, 128, (nn_, kk), dtype=torch.int8, generator=torch.Generator().manual_seed(i))
        ws.append((w, torch.rand(nn_) * 1e-3))
    if ops is not None:
        packed = [(ops.convert_weight_packed(w), s) for (w, s) in ws]
        return (lambda: [ops.int8_scaled_mm_with_quant(x, pw, s, None, dtype, True)
                         for (pw, s), x in zip(packed, xs)]), "int8_scaled_mm_with_quant"
    from oracle import quant_oracle as qo

    return (lambda: [qo.w8a8_int8_linear(x, w, s.view(-1, 1)) for (w, s), x in zip(ws, xs)]), "oracle port int8"


def run_reference_cpu(args, steps: int, warmup: int, bs: int, budget_s: float):
    """Times the reference CPU path.  A step runs `layers_per_step` layer passes (KV write + attention + the four
    linears) over ONE layer's tensors (>= 1 GB, far larger than the CPU caches); layers_per_step = all layers when the
    whole run fits `budget_s`, else fewer and the step time is scaled up (flagged `extrapolated`)."""
    from oracle import build_ref
    from sglang_npu_b200 import synth
    from sglang_npu_b200.srt.model_executor.layer_stack import CONFIGS

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    ops = build_ref.load()
    cfg = CONFIGS[args.model]
    hq, hkv, d = cfg.num_attention_heads, cfg.num_kv_heads, cfg.head_dim
    dtype = torch.bfloat16
    if args.workload == "decode":
        ctx = args.ctx
        pb = synth.make_paged_batch([ctx] * bs, PAGE, seed=1, shared_prefix=args.shared_prefix)
        k, v = _cpu_kv(pb.num_slots + PAGE, hkv, d)
        q = torch.randn(bs, hq, d).to(dtype)
        key, val = torch.randn(bs, hkv, d).to(dtype), torch.randn(bs, hkv, d).to(dtype)
        loc = torch.stack([pb.req_to_token[pb.req_pool_indices[i], ctx - 1] for i in range(bs)]).long()
        o = torch.empty_like(q)
        linears, lin_kind = _cpu_linears(ops, cfg, args.quant, bs, dtype)
        tokens = bs
        if ops is not None:
            kind = "reference"
            scratch = torch.zeros((bs, hq, 8, d + 1), dtype=torch.float32)

            def layer():
                ops.decode_attention_cpu(q, k, v, o, key, val, loc, scratch, pb.req_to_token, pb.req_pool_indices,
                                         pb.seq_lens, d ** -0.5, 0.0)
                linears()
            what = f"decode_attention_cpu (incl. KV write) + 4x {lin_kind}"
        else:  # host without AVX512-BF16/AMX: time the oracle port instead and say so
            kind = "port"
            from oracle import attention_oracle as ao

            def layer():
                ao.set_kv_buffer(k, v, loc, key, val)
                ao.decode_attention(q, k, v, pb.req_to_token, pb.req_pool_indices, pb.seq_lens, d ** -0.5)
                linears()
            what = f"oracle port: set_kv + decode attention + 4x {lin_kind}"
        desc = f"bs={bs}, ctx={ctx}"
    else:
        reqs, new, prefix = args.reqs, args.new, args.prefix
        pb = synth.make_paged_batch([prefix + new] * reqs, PAGE, seed=1, shared_prefix=prefix)
        k, v = _cpu_kv(pb.num_slots + PAGE, hkv, d)
        T = reqs * new
        q = torch.randn(T, hq, d).to(dtype)
        kx, vx = torch.randn(T, hkv, d).to(dtype), torch.randn(T, hkv, d).to(dtype)
        o = torch.empty_like(q)
        seq = torch.full((reqs,), prefix + new, dtype=torch.int64)
        ext = torch.full((reqs,), new, dtype=torch.int32)       # index dtype of req_to_token (extend.cpp:651-665)
        start = torch.arange(reqs, dtype=torch.int32) * new
        linears, lin_kind = _cpu_linears(ops, cfg, args.quant, T, dtype)
        tokens = T
        if ops is not None:
            kind = "reference"

            def layer():
                ops.extend_attention_cpu(q, kx, vx, o, k, v, pb.req_to_token, pb.req_pool_indices, seq, ext, start, new,
                                         d ** -0.5, 0.0)
                linears()
            what = f"extend_attention_cpu + 4x {lin_kind}"
        else:
            kind = "port"
            from oracle import attention_oracle as ao

            loc = torch.cat([pb.req_to_token[pb.req_pool_indices[i], prefix:prefix + new] for i in range(reqs)]).long()

            def layer():
                ao.set_kv_buffer(k, v, loc, kx, vx)
                ao.extend_attention(q, k, v, pb.req_to_token, pb.req_pool_indices, seq,
                                    torch.full((reqs,), prefix, dtype=torch.int64), ext, d ** -0.5)
                linears()
            what = f"oracle port: set_kv + extend attention + 4x {lin_kind}"
        desc = f"{reqs} x {new} new tokens over a {prefix}-token prefix"

    layer()  # one untimed pass: page in, size the sample
    t0 = time.perf_counter()
    layer()
    t_layer = time.perf_counter() - t0
    per_step = max(1, min(cfg.num_layers, int(budget_s / max(steps + warmup, 1) / max(t_layer, 1e-6))))
    for _ in range(warmup):
        for _ in range(per_step):
            layer()
    t0 = time.perf_counter()
    for _ in range(steps):
        for _ in range(per_step):
            layer()
    dt_step = (time.perf_counter() - t0) / steps
    step_s = dt_step * cfg.num_layers / per_step
    return {"value": tokens / step_s, "ms_per_layer": dt_step / per_step * 1e3, "ms_per_step": step_s * 1e3, "cores": cores,
            "kind": kind, "steps_timed": steps, "warmup_run": warmup, "layers_timed_per_step": per_step,
            "layers_per_step": cfg.num_layers, "extrapolated": per_step < cfg.num_layers,
            "sample": f"{per_step} of {cfg.num_layers} layer passes per step at full size ({desc}): {what}, every pass "
                      f"over one layer's tensors (larger than the CPU caches); step time = measured x "
                      f"{cfg.num_layers}/{per_step}; embedding / lm_head excluded; {steps} timed steps, {warmup} warm-up"}


# ---------------------------------------------------------------------------------------------------------
def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=DEFAULTS["workload"], choices=["decode", "prefill"])
    ap.add_argument("--model", default=DEFAULTS["model"], choices=sorted(MODEL_LABEL))
    ap.add_argument("--quant", default=DEFAULTS["quant"], choices=sorted(QUANT))
    ap.add_argument("--bs", type=int, default=DEFAULTS["bs"], help="decode batch per GPU (global with --fixed-global-batch)")
    ap.add_argument("--ctx", type=int, default=DEFAULTS["ctx"])
    ap.add_argument("--shared-prefix", type=int, default=0, dest="shared_prefix",
                    help="decode: leading tokens every request shares (radix-cache hit, config 5)")
    ap.add_argument("--fixed-global-batch", action="store_true", dest="fixed_global_batch",
                    help="decode: --bs is the GLOBAL batch at every N (strong scaling; configs 4 and 5 are quoted so)")
    ap.add_argument("--reqs", type=int, default=4, help="prefill: requests per chunk")
    ap.add_argument("--new", type=int, default=2048, help="prefill: new tokens per request")
    ap.add_argument("--prefix", type=int, default=6144, help="prefill: cached prefix tokens per request")
    ap.add_argument("--layers", type=int, default=0, help="debug only: fewer layers (marks the line invalid)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--timeline-out", default="", dest="timeline_out",
                    help="after the timed region, write rank 0's per-kernel device times of 3 graph replays (decode) / 2 steps (prefill) "
                         "(torch.profiler / CUPTI; works under torchrun where ncu does not) to this CSV")
    ap.add_argument("--kv-dtype", default=DEFAULTS["kv_dtype"], choices=["bf16", "fp8"], dest="kv_dtype",
                    help="fp8 = --kv-cache-dtype fp8_e4m3 variant")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    return args


def reference_arm(args):
    world = max(args.gpus, 1)
    bs = args.bs if (args.fixed_global_batch or args.workload == "prefill") else args.bs * world
    r = run_reference_cpu(args, args.steps, args.warmup, bs, budget_s=75.0)
    wl = (f"{args.model} {QUANT[args.quant][2]} decode, global batch {bs}, ctx {args.ctx}, page_size {PAGE}"
          if args.workload == "decode" else
          f"{args.model} {QUANT[args.quant][2]} prefill chunk {args.reqs} x {args.new} tokens, prefix {args.prefix}, "
          f"page_size {PAGE}")
    line = {"impl": "reference", "metric": metric_name(args), "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "strong" if (args.fixed_global_batch or args.workload == "prefill") else "weak",
            "vs_baseline": None, "dtype": QUANT[args.quant][3], "data": "synthetic",
            "config": {"workload": wl + ", reference CPU kernels (sgl-kernel/csrc/cpu) on host cores"},
            "steps_timed": r["steps_timed"], "layers_timed_per_step": r["layers_timed_per_step"],
            "layers_per_step": r["layers_per_step"], "extrapolated": r["extrapolated"],
            "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": r["kind"],
                             "sample": r["sample"]},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    if not is_headline(args):
        line["variant"] = "not the BASELINE headline configuration"
    print(json.dumps(line), flush=True)


def fill_kv_pool(pool, dev, kv_bytes: int):
    g = torch.Generator(device=dev).manual_seed(7)
    for buf in pool.k_buffer + pool.v_buffer:
        if kv_bytes == 2:
            buf.normal_(generator=g)  # N(0,1) like the reference tests
        else:  # the same distribution, stored as e4m3 bytes
            for lo in range(0, buf.shape[0], 1 << 16):
                part = buf[lo: lo + (1 << 16)]
                part.copy_(torch.empty(part.shape, dtype=torch.bfloat16, device=dev).normal_(generator=g)
                           .to(torch.float8_e4m3fn).view(torch.uint8))


def time_wrapped(modules_and_names, fn):
    """Run fn() with the named kernel wrappers bracketed by CUDA event pairs; returns {name: [ms, ...]}."""
    rec = {}
    saved = []
    for mod, name in modules_and_names:
        orig = getattr(mod, name)
        saved.append((mod, name, orig))

        def timed(*a, __orig=orig, __name=name, **kw):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            r = __orig(*a, **kw)
            e1.record()
            rec.setdefault(__name, []).append((e0, e1, a, kw))
            return r

        setattr(mod, name, timed)
    try:
        fn()
        torch.cuda.synchronize()
    finally:
        for mod, name, orig in saved:
            setattr(mod, name, orig)
    return rec


def kernel_timeline(fn, reps: int, out_path, barrier):
    """Per-kernel device time of `reps` calls of fn() on this rank (CUPTI through torch.profiler; kernels inside a
    replayed CUDA graph are recorded too).  Every rank runs the calls (collectives need all of them); only the rank
    given an out_path writes the CSV: kernel, launches, total_us, avg_us, share.  With programmatic dependent launch
    on, a kernel's CUPTI duration includes the time it spent waiting under its predecessor's tail (B200_PDL=0 gives
    each kernel's own time)."""
    from torch.profiler import ProfilerActivity, profile

    barrier()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(reps):
            fn()
        barrier()
    if out_path is None:
        return
    agg = {}
    for ev in prof.events():
        if getattr(ev, "device_type", None) is not None and str(ev.device_type).endswith("CUDA") and ev.device_time_total > 0:
            a = agg.setdefault(ev.name, [0, 0.0])
            a[0] += 1
            a[1] += ev.device_time_total
    tot = sum(a[1] for a in agg.values()) or 1.0
    with open(out_path, "w") as fh:
        fh.write(f"# {reps} graph replays, rank 0, device time per kernel (us); total {tot:.1f} us\n")
        fh.write("kernel,launches,total_us,avg_us,share\n")
        for name, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            fh.write(f"\"{name[:160]}\",{n},{us:.1f},{us / n:.2f},{us / tot:.4f}\n")


def tensor_peak_probe(kind: str, dev):
    """cuBLASLt 8192^3 through torch, best of 10 (the way MEASURED_PEAKS.json's bf16 number was taken)."""
    n = 8192
    try:
        if kind == "int8":
            a = torch.randint(-8, 8, (n, n), dtype=torch.int8, device=dev)
            b = torch.randint(-8, 8, (n, n), dtype=torch.int8, device=dev).t()
            f = lambda: torch._int_mm(a, b)  # noqa: E731
            what = "torch._int_mm int8 8192^3 (cuBLASLt)"
        else:
            a = torch.randn(n, n, device=dev).to(torch.float8_e4m3fn)
            b = torch.randn(n, n, device=dev).to(torch.float8_e4m3fn).t()
            one = torch.ones((), device=dev)
            f = lambda: torch._scaled_mm(a, b, scale_a=one, scale_b=one, out_dtype=torch.bfloat16)  # noqa: E731
            what = "torch._scaled_mm e4m3 8192^3 (cuBLASLt, tensor-wise scales)"
        for _ in range(3):
            f()
        best = 1e9
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            f()
            e1.record()
            torch.cuda.synchronize(dev)
            best = min(best, e0.elapsed_time(e1))
        return 2 * n ** 3 / (best * 1e-3) / 1e12, what
    except Exception as e:  # library refuses the shape / dtype on this box
        return None, f"probe failed: {e}"


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        if rank == 0:
            reference_arm(args)
        return 0

    import torch.distributed as dist

    from sglang_npu_b200 import kernels, synth
    from sglang_npu_b200.srt.distributed import parallel_state
    from sglang_npu_b200.srt.layers.quantization import get_quantization_config
    from sglang_npu_b200.srt.mem_cache.allocator import PagedTokenToKVPoolAllocator
    from sglang_npu_b200.srt.model_executor.layer_stack import CONFIGS
    from sglang_npu_b200.srt.model_executor.model_runner import HostDecodeBatch, HostExtendBatch, HotPathRunner

    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: the B200 hot path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        parallel_state.init_tensor_parallel()
    tp = world
    cfg = CONFIGS[args.model]
    n_layers = args.layers or cfg.num_layers
    qname, act_dtype, qlabel, arith = QUANT[args.quant]
    qc = get_quantization_config(qname).from_config(QCFG)
    kv_bytes = 1 if args.kv_dtype == "fp8" else 2
    kv_dtype = torch.float8_e4m3fn if kv_bytes == 1 else None
    peaks, peak_kind = measured_peaks()
    hkv_local, hq_local = cfg.get_num_kv_heads(tp), cfg.num_attention_heads // tp

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def max_over_ranks(x: float) -> float:
        t = torch.tensor([x], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    line_extra = {}
    if args.workload == "decode":
        bs = args.bs if args.fixed_global_batch else args.bs * tp  # weak scaling by default: per-GPU KV bytes fixed
        ctx = args.ctx
        total_e2e = args.warmup + args.steps
        headroom = min((total_e2e + 8 + PAGE - 1) // PAGE * PAGE, 1024)  # decode-loop steps before the loop state resets
        new_pages = bs * (headroom // PAGE + 1)
        pb = synth.make_paged_batch([ctx] * bs, PAGE, seed=1, width=ctx + headroom + 4, max_reqs=bs + 3,
                                    shared_prefix=args.shared_prefix, extra_pages=new_pages + 3)
        runner = HotPathRunner(cfg, qc, max_reqs=pb.req_to_token.shape[0], max_total_tokens=pb.num_slots, page_size=PAGE,
                               context_len=ctx + headroom, device=dev, num_layers=n_layers, tp_size=tp, seed=1234,
                               kv_cache_dtype=kv_dtype, dtype=act_dtype)
        runner.req_to_token_pool.req_to_token.copy_(pb.req_to_token)
        fill_kv_pool(runner.token_to_kv_pool, dev, kv_bytes)
        used = torch.unique(pb.req_to_token // PAGE)
        free0 = torch.tensor(sorted(set(range(1, pb.num_slots // PAGE + 1)) - set(used.tolist())), dtype=torch.int64,
                             device=dev)
        allocator = PagedTokenToKVPoolAllocator(pb.num_slots, PAGE, act_dtype, dev)
        allocator.free_pages = free0.clone()
        runner.attach_allocator(allocator)

        out_loc = torch.stack([pb.req_to_token[pb.req_pool_indices[i], ctx - 1] for i in range(bs)]).long()
        host = HostDecodeBatch(input_ids=torch.randint(0, cfg.vocab_size, (bs,), dtype=torch.int64).pin_memory(),
                               req_pool_indices=pb.req_pool_indices.clone().pin_memory(),
                               seq_lens=pb.seq_lens.clone().pin_memory(), out_cache_loc=out_loc.pin_memory())
        d_in = {k: getattr(host, k).to(dev) for k in ("input_ids", "req_pool_indices", "seq_lens", "out_cache_loc")}

        kernels.reset_launch_count()
        runner.capture_decode_graph(bs, shared_prefix_len=args.shared_prefix)
        launches_per_step = kernels.launch_count() // 3  # 2 warm-up passes + 1 captured pass

        # ---- (1) device-resident steps: `value` ----------------------------------------------------------
        for _ in range(args.warmup):
            runner.replay_decode(**d_in)
        sampler = ClockSampler(local_rank)
        barrier()
        if rank == 0:
            sampler.start()
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
        evs[0].record()
        for i in range(args.steps):
            runner.replay_decode(**d_in)
            evs[i + 1].record()
        barrier()
        clocks = sampler.stop() if rank == 0 else None
        ms_per_step = max_over_ranks(evs[0].elapsed_time(evs[-1])) / args.steps
        per = sorted(evs[i].elapsed_time(evs[i + 1]) for i in range(args.steps))
        ms_median = max_over_ranks(per[len(per) // 2])
        tokens_per_step = bs
        value = bs / (ms_per_step * 1e-3)

        # ---- (2) end to end: a decode loop through decode_loop_step(), pinned host in / out -----------------
        def reset_loop():
            allocator.free_pages = free0.clone()
            host.seq_lens.copy_(pb.seq_lens)

        done = 0
        for _ in range(args.warmup):
            host.seq_lens.add_(1)
            host.input_ids.copy_(runner.decode_loop_step(host))
            done += 1
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            if done == headroom:
                reset_loop()
                done = 0
            host.seq_lens.add_(1)  # the scheduler's prepare_for_decode: seq_lens += 1 (schedule_batch.py:1560)
            ids = runner.decode_loop_step(host)
            host.input_ids.copy_(ids)  # the sampled ids are the next step's inputs
            done += 1
        barrier()
        e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1e3)
        e2e_value = bs / (e2e_ms / args.steps * 1e-3)
        h2d = sum(t.numel() * t.element_size() for t in (host.input_ids, host.req_pool_indices, host.seq_lens))
        d2h = bs * 8 + 8  # sampled ids + the allocator's page counter
        reset_loop()

        if args.timeline_out:
            kernel_timeline(lambda: runner.replay_decode(**d_in), 3, args.timeline_out if rank == 0 else None, barrier)

        # ---- (3) roofline of the dominant kernel, measured live with CUDA events on the launch stream --------
        import sglang_npu_b200.srt.layers.attention.b200_backend as bmod

        name = "decode_attention_shared_prefix_fwd" if runner.attn_backend.forward_metadata.shared_prefix_len else \
            "decode_attention_fwd"

        def eager_passes():
            for _ in range(3):  # seq_lens_cpu=None: the same (worst-case-planned) chunking as the captured graph
                runner.forward_decode(d_in["input_ids"], d_in["req_pool_indices"], d_in["seq_lens"], d_in["out_cache_loc"],
                                      seq_lens_cpu=None, shared_prefix_len=args.shared_prefix)

        rec = time_wrapped([(bmod.kernels, name)], eager_passes)
        pairs = rec[name]
        durs = sorted(a.elapsed_time(b) for a, b, _, _ in pairs[len(pairs) // 3:])  # drop the first pass
        attn_ms_eager = durs[len(durs) // 2]  # median over every layer of two passes

        # The same launch group the way the step runs it — from a CUDA graph, launches chained by programmatic
        # dependent launch: every layer's attention (its own KV buffers: 1.07 GB per launch, L2-cold) captured back to
        # back in one graph, CUDA events around each of three replays.  The eager figure above carries the host launch
        # gap between the split-KV kernel and its merge; it is kept as `ms_per_launch_eager`.
        def attn_graph_ms():
            from sglang_npu_b200.srt.model_executor.forward_batch_info import ForwardMode

            fb = runner._make_batch(ForwardMode.DECODE, d_in["input_ids"], d_in["req_pool_indices"], d_in["seq_lens"],
                                    d_in["out_cache_loc"])
            layers = [lyr for lyr in runner.model.layers]
            q = torch.randn((bs, hq_local * cfg.head_dim), device=dev).to(act_dtype)

            def run():
                runner.attn_backend.init_forward_metadata_capture_cuda_graph(bs, bs, d_in["req_pool_indices"],
                                                                            d_in["seq_lens"], None, ForwardMode.DECODE, None)
                for lyr in layers:
                    lyr.attn(q, None, None, fb, save_kv_cache=False, int8_out=bool(getattr(lyr, "fused_int8", False) and getattr(lyr, "fuse_attn_quant", False)),
                             kv_in_cache=True)

            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                run()
            torch.cuda.current_stream(dev).wait_stream(side)
            torch.cuda.synchronize(dev)
            g_att = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g_att):
                run()
            g_att.replay()
            torch.cuda.synchronize(dev)
            reps = []
            for _ in range(3):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                g_att.replay()
                e1.record()
                torch.cuda.synchronize(dev)
                reps.append(e0.elapsed_time(e1) / len(layers))
            return sorted(reps)[1]

        try:
            attn_ms = attn_graph_ms()
            attn_how = ("CUDA events around graph replays of every layer's attention launch group back to back (median of 3 "
                        "replays / layers); ms_per_launch_eager = the same group timed launch by launch in eager passes")
        except Exception as e:  # keep the line: fall back to the eager figure and say so
            attn_ms = attn_ms_eager
            attn_how = f"eager passes, CUDA events per launch group (graph timing failed: {str(e)[:80]})"
        # algorithmic bytes per launch: KV (unique bytes when a prefix is shared) + q/o + page ids
        sp = args.shared_prefix
        kv_tokens = sp + bs * (ctx - sp)
        attn_bytes = (kv_tokens * 2 * hkv_local * cfg.head_dim * kv_bytes + bs * hq_local * cfg.head_dim * 2 * 2
                      + 4 * bs * (ctx // PAGE))
        achieved = attn_bytes / (attn_ms * 1e-3) / 1e9
        kname = "decode_attn_fp8kv_kernel" if kv_bytes == 1 else "decode_attn_kernel"
        roofline = {"kernel": (("shared-prefix pass (tcgen05) + " if sp else "") + kname +
                               " (+ decode_merge_quant_kernel: LSE merge fused with the int8 quant of the o_proj input), "
                               "one launch group per layer"),
                    "bound": "hbm", "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                    "frac": achieved / peaks["hbm_gbs"],
                    "peak_kind": f"{peak_kind} (MEASURED_PEAKS.json hbm_gbs, copy bandwidth)",
                    "frac_of_nominal_8TBs": achieved / 8000.0, "bytes_per_launch": attn_bytes, "ms_per_launch": attn_ms,
                    "ms_per_launch_eager": attn_ms_eager, "timing": attn_how,
                    "share_of_step": attn_ms * n_layers / ms_per_step, "traffic": None}
        if sp:
            roofline["bytes_note"] = "UNIQUE KV bytes (the shared prefix counted once); logical bytes = " \
                                     f"{bs * ctx * 2 * hkv_local * cfg.head_dim * kv_bytes}"
        traffic_file = os.path.join(ROOT, "profiles", "decode_attn_traffic.json")
        if os.path.exists(traffic_file) and is_headline(args) and tp == 1:
            try:
                roofline["traffic"] = json.load(open(traffic_file)).get("dram_bytes_per_launch")
                roofline["traffic_source"] = "profiles/decode_attn_traffic.json (ncu --set full capture of this kernel " \
                                             "at this configuration, not re-measured in this run)"
            except Exception:
                pass
        workload = (f"{args.model} {qlabel} decode step (embedding + {n_layers} layers + {str(act_dtype)[6:]} lm_head + "
                    f"greedy argmax), global batch {bs}, ctx {ctx}"
                    + (f", shared prefix {sp}" if sp else "") + f", page_size {PAGE}, "
                    f"{'bf16' if kv_bytes == 2 and act_dtype == torch.bfloat16 else 'f16' if kv_bytes == 2 else 'fp8_e4m3'} KV")
        weights_gb = sum(k * n for k, n in linear_shapes(cfg, tp)) * (0.5 if args.quant == "awq" else 1.0) / 1e9
        l2 = (f"inputs larger than L2: {attn_bytes / 1e9:.2f} GB of KV + {weights_gb:.2f} GB of weights per layer vs "
              "126 MB L2")
        scaling = "strong" if args.fixed_global_batch else "weak"
        e2e_note = ("decode loop: seq_lens advance by 1 per step from ctx; KV slot allocated on device by the fused "
                    "alloc_decode launch")
        global_batch, seq_len = bs, ctx
    else:
        # ---- prefill: one chunked-prefill step ----------------------------------------------------------------
        reqs, new, prefix = args.reqs, args.new, args.prefix
        T = reqs * new
        pb = synth.make_paged_batch([prefix + new] * reqs, PAGE, seed=1, width=prefix + new + 4, max_reqs=reqs + 3,
                                    shared_prefix=prefix)
        runner = HotPathRunner(cfg, qc, max_reqs=pb.req_to_token.shape[0], max_total_tokens=pb.num_slots, page_size=PAGE,
                               context_len=prefix + new, device=dev, num_layers=n_layers, tp_size=tp, seed=1234,
                               kv_cache_dtype=kv_dtype, dtype=act_dtype, tp_comm_tokens=T)
        runner.req_to_token_pool.req_to_token.copy_(pb.req_to_token)
        fill_kv_pool(runner.token_to_kv_pool, dev, kv_bytes)  # the cached prefix
        rpi = pb.req_pool_indices
        host = HostExtendBatch(
            input_ids=torch.randint(0, cfg.vocab_size, (T,), dtype=torch.int64).pin_memory(),
            req_pool_indices=rpi.clone().pin_memory(),
            seq_lens=torch.full((reqs,), prefix + new, dtype=torch.int64).pin_memory(),
            prefix_lens=torch.full((reqs,), prefix, dtype=torch.int64).pin_memory(),
            out_cache_loc=torch.cat([pb.req_to_token[rpi[i], prefix:prefix + new] for i in range(reqs)]).long().pin_memory(),
            positions=torch.arange(prefix, prefix + new, dtype=torch.int64).repeat(reqs).pin_memory())
        d_in = {k: getattr(host, k).to(dev) for k in
                ("input_ids", "req_pool_indices", "seq_lens", "prefix_lens", "out_cache_loc", "positions")}
        ext_cpu, pre_cpu = [new] * reqs, [prefix] * reqs

        def step():
            return runner.forward_extend(d_in["input_ids"], d_in["req_pool_indices"], d_in["seq_lens"], d_in["prefix_lens"],
                                         d_in["out_cache_loc"], d_in["positions"], seq_lens_cpu=host.seq_lens,
                                         extend_seq_lens_cpu=ext_cpu, extend_prefix_lens_cpu=pre_cpu, greedy=True)

        step()
        torch.cuda.synchronize(dev)
        kernels.reset_launch_count()
        step()
        launches_per_step = kernels.launch_count()
        for _ in range(args.warmup):
            step()
        sampler = ClockSampler(local_rank)
        barrier()
        if rank == 0:
            sampler.start()
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
        evs[0].record()
        for i in range(args.steps):
            step()
            evs[i + 1].record()
        barrier()
        clocks = sampler.stop() if rank == 0 else None
        ms_per_step = max_over_ranks(evs[0].elapsed_time(evs[-1])) / args.steps
        per = sorted(evs[i].elapsed_time(evs[i + 1]) for i in range(args.steps))
        ms_median = max_over_ranks(per[len(per) // 2])
        tokens_per_step = T
        value = T / (ms_per_step * 1e-3)

        for _ in range(args.warmup):
            runner.prefill_step(host)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            runner.prefill_step(host)
        barrier()
        e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1e3)
        e2e_value = T / (e2e_ms / args.steps * 1e-3)
        h2d, d2h = host.nbytes(), reqs * 8

        if args.timeline_out:
            kernel_timeline(step, 2, args.timeline_out if rank == 0 else None, barrier)

        # ---- roofline: the four GEMMs of a layer, timed live; the largest (gate_up) is the dominant kernel -------
        gemm_fn = {"int8": "int8_scaled_mm", "fp8": "fp8_blockwise_scaled_mm", "fp8_chan": "fp8_scaled_mm",
                   "awq": "awq_gemm"}[args.quant]
        import sglang_npu_b200.srt.layers.attention.b200_backend as bmod

        rec = time_wrapped([(kernels, gemm_fn), (bmod.kernels, "extend_attention_fwd")],
                           lambda: [step() for _ in range(3)])
        shapes = linear_shapes(cfg, tp)
        calls = rec[gemm_fn][len(rec[gemm_fn]) // 3:]
        by_shape = {}
        for e0, e1, a, kw in calls:
            if args.quant == "awq":
                kk, nn_ = a[1].shape[0], a[1].shape[1] * 8
            else:
                kk, nn_ = a[1].shape[0], a[1].shape[1]
            by_shape.setdefault((kk, nn_), []).append(e0.elapsed_time(e1))
        gemms = []
        for (kk, nn_), ds in by_shape.items():
            ms = sorted(ds)[len(ds) // 2]  # median: a power-cap clock dip during one pass must not skew the figure
            gemms.append({"k": kk, "n": nn_, "ms": ms, "tflops": 2 * T * kk * nn_ / (ms * 1e-3) / 1e12})
        gemms.sort(key=lambda g_: -g_["k"] * g_["n"])
        dom = gemms[0]
        tot_ms = sum(g_["ms"] for g_ in gemms)
        tot_flops = sum(2 * T * g_["k"] * g_["n"] for g_ in gemms)
        att = rec.get("extend_attention_fwd", [])
        att = att[len(att) // 3:]
        att_d = sorted(a.elapsed_time(b) for a, b, _, _ in att)
        attn_ms = att_d[len(att_d) // 2] if att_d else 0.0
        attn_flops = 4 * cfg.head_dim * hq_local * reqs * (new * prefix + new * (new + 1) // 2)
        probe, probe_what = tensor_peak_probe("int8" if args.quant == "int8" else "fp8", dev) \
            if args.quant != "awq" else (peaks["bf16_tflops"], "MEASURED_PEAKS.json bf16_tflops (f16 operands after dequant)")
        peak = probe or 4500.0
        roofline = {"kernel": f"{gemm_fn} gate_up GEMM [M={T}, K={dom['k']}, N={dom['n']}] (tcgen05, persistent)",
                    "bound": "tensor", "achieved": dom["tflops"], "peak": peak, "unit": "TFLOP/s",
                    "frac": dom["tflops"] / peak,
                    "peak_kind": (f"measured in this run: {probe_what}" if probe else
                                  f"nominal dense 8-bit peak ({probe_what})") +
                                 "; MEASURED_PEAKS.json holds no 8-bit tensor peak",
                    "frac_of_nominal_4500": dom["tflops"] / 4500.0,
                    "frac_of_2x_measured_bf16": dom["tflops"] / (2 * peaks["bf16_tflops"]),
                    "flops_per_launch": 2 * T * dom["k"] * dom["n"], "ms_per_launch": dom["ms"],
                    "share_of_step": dom["ms"] * n_layers / ms_per_step, "traffic": None,
                    "all_gemms": gemms, "all_gemms_tflops": tot_flops / (tot_ms * 1e-3) / 1e12,
                    "all_gemms_share_of_step": tot_ms * n_layers / ms_per_step,
                    "extend_attention": {"ms_per_launch": attn_ms, "tflops": attn_flops / (attn_ms * 1e-3) / 1e12 if attn_ms else None,
                                         "frac_of_bf16_peak": (attn_flops / (attn_ms * 1e-3) / 1e12 / peaks["bf16_tflops"]) if attn_ms else None,
                                         "share_of_step": attn_ms * n_layers / ms_per_step}}
        workload = (f"{args.model} {qlabel} chunked-prefill step (embedding + {n_layers} layers + lm_head of the last "
                    f"tokens + greedy argmax), {reqs} requests x {new} new tokens = {T}-token chunk over a "
                    f"{prefix}-token radix-cache prefix per request, page_size {PAGE}, "
                    f"{'bf16' if kv_bytes == 2 else 'fp8_e4m3'} KV")
        l2 = f"inputs larger than L2: {sum(k * n for k, n in shapes) / 1e6:.0f} MB of weights + " \
             f"{T * cfg.hidden_size * 2 / 1e6:.0f} MB activations per layer vs 126 MB L2"
        scaling = "strong"
        e2e_note = "prefill_step(): H2D of ids / locs / positions, eager launches, D2H of the sampled ids"
        global_batch, seq_len = reqs, prefix + new

    if rank != 0:
        _finish(world, dev)
        return 0

    cpu_baseline = None
    if world == 1 and not args.no_cpu_baseline:
        r = run_reference_cpu(args, 3, 1, global_batch if args.workload == "decode" else 0, budget_s=20.0)
        cpu_baseline = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": r["kind"], "sample": r["sample"],
                        "ms_per_layer": r["ms_per_layer"], "extrapolated": r["extrapolated"]}

    comm = getattr(runner, "tp_comm", None)
    line = {"metric": metric_name(args), "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "ms_per_step_median": ms_median, "higher_is_better": True,
            "scaling": scaling, "vs_baseline": None, "dtype": arith, "data": "synthetic",
            "config": {"workload": workload, "global_batch": global_batch, "seq_len": seq_len, "parallelism": f"tp{tp}",
                       "cuda_graph": args.workload == "decode",
                       "tp_exchange": ("none (tp1)" if tp == 1 else
                                       "fused reduce + rmsnorm (+ int8 quant) over NVLink peer memory (csrc/tp_comm.cu); "
                                       "vocab-parallel argmax (8 B/row NCCL max all-reduce)"
                                       if comm is not None else "NCCL all-reduce"),
                       "l2_policy": l2, "tok_s_per_gpu": value / world, "tokens_per_step": tokens_per_step},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "what": e2e_note},
            "gpu_launches": launches_per_step * args.steps, "gpu_launches_per_step": launches_per_step,
            "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu_baseline}
    line.update(line_extra)
    if args.layers:
        line["invalid"] = f"debug run with {n_layers} layers"
    if not is_headline(args):
        line["variant"] = "not the BASELINE headline configuration (configs[1]); reported next to it"
    print(json.dumps(line), flush=True)
    _finish(world, dev)
    return 0


def _finish(world, dev):
    """Leave without tearing NCCL down: destroy_process_group() can block for minutes when captured CUDA graphs
    still hold the communicator; the measurements are done, so synchronise, meet the other ranks and exit hard."""
    if world <= 1:
        return
    import torch.distributed as dist

    torch.cuda.synchronize(dev)
    try:
        dist.barrier()
        torch.cuda.synchronize(dev)
    finally:
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


if __name__ == "__main__":
    sys.exit(main())
