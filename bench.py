#!/usr/bin/env python
"""bench.py — decode tok/s for Llama-3.1-8B W8A8-int8, bs=64/GPU, 4k ctx, page_size=16 (BASELINE.json
configs[1]) through the B200 hot path, with the roofline of the dominant kernel and the reference's CPU
path timed beside it.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

* a "step" = one decode forward of the whole layer stack (32 layers + lm_head + greedy argmax) over one
  synthetic batch; CUDA-graph replay, inputs resident in HBM -> `value` (whole-job tok/s).
* `e2e` = the same step through the public call HotPathRunner.decode_step() with PINNED HOST index tensors:
  per step H2D of input_ids / req_pool_indices / seq_lens / out_cache_loc, D2H of the sampled token ids,
  and the next step's input_ids are the ids that came back (a real decode loop).
* N > 1: tensor parallel over N GPUs of one box (heads + column/row-parallel linears sharded, NCCL
  all-reduce after o_proj/down_proj), weak scaling: global batch = 64 * N so per-GPU KV bytes stay fixed.
* `--impl reference`: the reference's own CPU implementation of the path (oracle/_ref: decode_attention_cpu +
  int8_scaled_mm_with_quant compiled from /root/reference) on the host cores, one layer per step, scaled
  to the same metric.
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

METRIC = "decode tok/s (Llama-3.1-8B W8A8-int8, bs64/GPU, 4k ctx, page16)"
UNIT = "tok/s"
MODEL = "llama-3.1-8b"
BS_PER_GPU = 64
CTX = 4096
PAGE = 16


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


# ---------------------------------------------------------------------------------------------------------
# reference arm / CPU baseline: the reference's own CPU kernels on one layer of the same workload
# ---------------------------------------------------------------------------------------------------------
def cpu_layer_workload(bs: int, ctx: int):
    """Host tensors for ONE layer of the workload (KV pool, page table, int8 weights)."""
    from sglang_npu_b200 import synth
    from sglang_npu_b200.srt.model_executor.layer_stack import CONFIGS

    cfg = CONFIGS[MODEL]
    pb = synth.make_paged_batch([ctx] * bs, PAGE, seed=1)
    slots = pb.num_slots + PAGE
    blk = torch.randn(1 << 22, generator=torch.Generator().manual_seed(0)).to(torch.bfloat16)
    n = slots * cfg.num_kv_heads * cfg.head_dim
    k = blk.repeat((n + blk.numel() - 1) // blk.numel())[:n].view(slots, cfg.num_kv_heads, cfg.head_dim).contiguous()
    v = k.flip(0).contiguous()
    shapes = [(cfg.hidden_size, (cfg.num_attention_heads + 2 * cfg.num_kv_heads) * cfg.head_dim),
              (cfg.num_attention_heads * cfg.head_dim, cfg.hidden_size),
              (cfg.hidden_size, 2 * cfg.intermediate_size), (cfg.intermediate_size, cfg.hidden_size)]
    weights = []
    for i, (kk, nn_) in enumerate(shapes):
        w = torch.randint(-127, 128, (nn_, kk), dtype=torch.int8, generator=torch.Generator().manual_seed(i))
        s = torch.rand(nn_) * 1e-3
        weights.append((w, s, kk, nn_))
    return cfg, pb, k, v, weights


def run_reference_cpu(steps: int, warmup: int, bs: int, ctx: int):
    """Times the reference CPU path: per step ONE layer (KV write + decode attention + the four W8A8 linears)."""
    from oracle import build_ref

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    ops = build_ref.load()
    cfg, pb, k, v, weights = cpu_layer_workload(bs, ctx)
    hq, hkv, d = cfg.num_attention_heads, cfg.num_kv_heads, cfg.head_dim
    q = torch.randn(bs, hq, d).to(torch.bfloat16)
    key = torch.randn(bs, hkv, d).to(torch.bfloat16)
    val = torch.randn(bs, hkv, d).to(torch.bfloat16)
    loc = torch.stack([pb.req_to_token[pb.req_pool_indices[i], ctx - 1] for i in range(bs)]).long()
    o = torch.empty_like(q)
    if ops is not None:
        kind = "reference"
        packed = [(ops.convert_weight_packed(w), s, kk, nn_) for (w, s, kk, nn_) in weights]
        logits = torch.zeros((bs, hq, 8, d + 1), dtype=torch.float32)
        xs = [torch.randn(bs, kk).to(torch.bfloat16) * 0.1 for (_, _, kk, _) in weights]

        def layer():
            ops.decode_attention_cpu(q, k, v, o, key, val, loc, logits, pb.req_to_token, pb.req_pool_indices,
                                     pb.seq_lens, d ** -0.5, 0.0)
            for (pw, s, kk, nn_), x in zip(packed, xs):
                ops.int8_scaled_mm_with_quant(x, pw, s, None, torch.bfloat16, True)
    else:  # host without AVX512-BF16/AMX: time the oracle port instead and say so
        kind = "port"
        from oracle import attention_oracle as ao
        from oracle import quant_oracle as qo

        xs = [torch.randn(bs, kk).to(torch.bfloat16) * 0.1 for (_, _, kk, _) in weights]

        def layer():
            ao.set_kv_buffer(k, v, loc, key, val)
            ao.decode_attention(q, k, v, pb.req_to_token, pb.req_pool_indices, pb.seq_lens, d ** -0.5)
            for (w, s, kk, nn_), x in zip(weights, xs):
                qo.w8a8_int8_linear(x, w, s.view(-1, 1))

    for _ in range(warmup):
        layer()
    t0 = time.perf_counter()
    for _ in range(steps):
        layer()
    dt = (time.perf_counter() - t0) / steps
    step_s = dt * cfg.num_layers  # one decode step = num_layers layers (lm_head / embedding not included)
    return {"value": bs / step_s, "ms_per_layer": dt * 1e3, "ms_per_step": step_s * 1e3, "cores": cores, "kind": kind,
            "sample": f"1 of {cfg.num_layers} layers per step at full batch (bs={bs}, ctx={ctx}): set_kv + "
                      f"decode_attention_cpu + 4x int8_scaled_mm_with_quant; step time = {cfg.num_layers} x layer "
                      f"time; lm_head/embedding excluded; {steps} timed layers"}


# ---------------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--bs", type=int, default=BS_PER_GPU, help="batch per GPU (default: the BASELINE config)")
    ap.add_argument("--ctx", type=int, default=CTX)
    ap.add_argument("--layers", type=int, default=0, help="debug only: fewer layers (marks the line invalid)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--kv-dtype", default="bf16", choices=["bf16", "fp8"],
                    help="fp8 = --kv-cache-dtype fp8_e4m3 variant (NOT the BASELINE config: the line is marked as a variant)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        if rank != 0:
            return 0
        bs = args.bs * max(args.gpus, 1)
        r = run_reference_cpu(max(min(args.steps, 8), 2), 1, bs, args.ctx)
        line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "int8", "data": "synthetic",
                "config": {"workload": f"{MODEL} W8A8-int8 decode, global batch {bs}, ctx {args.ctx}, page_size {PAGE}, "
                                       "reference CPU kernels (sgl-kernel/csrc/cpu) on host cores"},
                "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": r["kind"],
                                 "sample": r["sample"]},
                "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line), flush=True)
        return 0

    # ---- B200 arm -----------------------------------------------------------------------------------------
    import torch.distributed as dist

    from sglang_npu_b200 import kernels
    from sglang_npu_b200.srt.distributed import parallel_state
    from sglang_npu_b200.srt.layers.quantization.w8a8_int8 import B200W8A8Int8Config
    from sglang_npu_b200.srt.model_executor.layer_stack import CONFIGS
    from sglang_npu_b200.srt.model_executor.model_runner import HostDecodeBatch, HotPathRunner
    from sglang_npu_b200 import synth

    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: the B200 hot path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        parallel_state.init_tensor_parallel()
    tp = world
    bs = args.bs * tp  # weak scaling: per-GPU KV bytes fixed (Hkv / tp heads x bs * tp requests)
    ctx = args.ctx
    cfg = CONFIGS[MODEL]
    n_layers = args.layers or cfg.num_layers

    pb = synth.make_paged_batch([ctx] * bs, PAGE, seed=1, width=ctx + 4, max_reqs=bs + 3)
    runner = HotPathRunner(cfg, B200W8A8Int8Config(), max_reqs=pb.req_to_token.shape[0], max_total_tokens=pb.num_slots,
                           page_size=PAGE, context_len=ctx, device=dev, num_layers=n_layers, tp_size=tp, seed=1234,
                           kv_cache_dtype=torch.float8_e4m3fn if args.kv_dtype == "fp8" else None)
    kv_bytes = 1 if args.kv_dtype == "fp8" else 2
    runner.req_to_token_pool.req_to_token.copy_(pb.req_to_token)
    g = torch.Generator(device=dev).manual_seed(7)
    for buf in runner.token_to_kv_pool.k_buffer + runner.token_to_kv_pool.v_buffer:
        if kv_bytes == 2:
            buf.normal_(generator=g)  # KV of the 4k-token contexts, N(0,1) like the reference tests
        else:  # the same distribution, stored as e4m3 bytes (a random finite e4m3 pattern per N(0,1) draw)
            for lo in range(0, buf.shape[0], 1 << 16):
                part = buf[lo: lo + (1 << 16)]
                part.copy_(torch.empty(part.shape, dtype=torch.bfloat16, device=dev).normal_(generator=g)
                           .to(torch.float8_e4m3fn).view(torch.uint8))

    out_loc = torch.stack([pb.req_to_token[pb.req_pool_indices[i], ctx - 1] for i in range(bs)]).long()
    host = HostDecodeBatch(input_ids=torch.randint(0, cfg.vocab_size, (bs,), dtype=torch.int64).pin_memory(),
                           req_pool_indices=pb.req_pool_indices.clone().pin_memory(),
                           seq_lens=pb.seq_lens.clone().pin_memory(), out_cache_loc=out_loc.pin_memory())
    d_in = {k: getattr(host, k).to(dev) for k in ("input_ids", "req_pool_indices", "seq_lens", "out_cache_loc")}

    kernels.reset_launch_count()
    runner.capture_decode_graph(bs)
    launches_per_step = kernels.launch_count() // 3  # 2 warm-up passes + 1 captured pass

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---- (1) device-resident steps: `value` --------------------------------------------------------------
    for _ in range(args.warmup):
        runner.replay_decode(**d_in)
    sampler = ClockSampler(local_rank)
    barrier()
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        runner.replay_decode(**d_in)
    ev1.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms = ev0.elapsed_time(ev1)
    t = torch.tensor([ms], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_per_step = t.item() / args.steps
    value = bs / (ms_per_step * 1e-3)

    # ---- (2) end to end through decode_step(): pinned host in, pinned host out ------------------------------
    for _ in range(args.warmup):
        runner.decode_step(host)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ids = runner.decode_step(host)
        host.input_ids.copy_(ids)  # feed the sampled ids back, as a decode loop does
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    t = torch.tensor([e2e_ms], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = bs / (t.item() / args.steps * 1e-3)

    # ---- (3) roofline of the dominant kernel, measured live with CUDA events on the launch stream ------------
    #      (eager pass: every decode-attention launch of the layer loop is bracketed by an event pair)
    pairs = []
    orig = kernels.decode_attention_fwd

    def timed(*a, **kw):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        r = orig(*a, **kw)  # stage 1 + stage 2 (the LSE merge, fused with the int8 quant of the o_proj input)
        e1.record()
        pairs.append((e0, e1))
        return r

    kernels.decode_attention_fwd = timed
    import sglang_npu_b200.srt.layers.attention.b200_backend as bmod
    bmod.kernels.decode_attention_fwd = timed
    for _ in range(3):
        runner.forward_decode(d_in["input_ids"], d_in["req_pool_indices"], d_in["seq_lens"], d_in["out_cache_loc"],
                              seq_lens_cpu=None)  # same (worst-case-planned) chunking as the captured graph
    torch.cuda.synchronize(dev)
    kernels.decode_attention_fwd = orig
    bmod.kernels.decode_attention_fwd = orig
    durs = sorted(a.elapsed_time(b) for a, b in pairs[len(pairs) // 3:])  # drop the first pass
    attn_ms = sum(durs) / len(durs)
    hkv_local = cfg.get_num_kv_heads(tp)
    hq_local = cfg.num_attention_heads // tp
    # algorithmic bytes per launch: KV + q/o + page ids
    attn_bytes = bs * ctx * 2 * hkv_local * cfg.head_dim * kv_bytes + bs * hq_local * cfg.head_dim * 2 * 2 + 4 * bs * (ctx // PAGE)
    peaks, peak_kind = measured_peaks()
    achieved = attn_bytes / (attn_ms * 1e-3) / 1e9
    roofline = {"kernel": ("decode_attn_fp8kv_kernel" if kv_bytes == 1 else "decode_attn_kernel") + " (+ decode_merge_quant_kernel: LSE merge fused with the int8 quant of the o_proj input), one launch pair per layer", "bound": "hbm",
                "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": achieved / peaks["hbm_gbs"],
                "peak_kind": f"{peak_kind} (MEASURED_PEAKS.json hbm_gbs, copy bandwidth)",
                "frac_of_nominal_8TBs": achieved / 8000.0, "bytes_per_launch": attn_bytes, "ms_per_launch": attn_ms,
                "share_of_step": attn_ms * n_layers / ms_per_step, "traffic": None}
    traffic_file = os.path.join(ROOT, "profiles", "decode_attn_traffic.json")
    if os.path.exists(traffic_file) and kv_bytes == 2:
        try:
            roofline["traffic"] = json.load(open(traffic_file)).get("dram_bytes_per_launch")
        except Exception:
            pass

    if rank != 0:
        _finish(world, dev)
        return 0

    cpu_baseline = None
    if world == 1 and not args.no_cpu_baseline:
        r = run_reference_cpu(6, 1, bs, ctx)
        cpu_baseline = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": r["kind"], "sample": r["sample"],
                        "ms_per_layer": r["ms_per_layer"]}

    h2d = host.nbytes()
    d2h = bs * 8
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int8",
            "data": "synthetic",
            "config": {"workload": f"{MODEL} W8A8-int8 decode step ({n_layers} layers + bf16 lm_head + greedy argmax), "
                                   f"global batch {bs}, ctx {ctx}, page_size {PAGE}, {'bf16' if kv_bytes == 2 else 'fp8_e4m3'} KV",
                       "global_batch": bs, "seq_len": ctx, "parallelism": f"tp{tp}", "cuda_graph": True,
                       "tp_exchange": ("none (tp1)" if tp == 1 else
                                       "fused reduce+rmsnorm+int8-quant over NVLink peer memory (csrc/tp_comm.cu)"
                                       if getattr(runner, "tp_comm", None) is not None else "NCCL all-reduce"),
                       "l2_policy": "inputs larger than L2: 1.07 GB of KV + 0.22 GB of weights per layer vs 126 MB L2",
                       "tok_s_per_gpu": value / world},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": launches_per_step * args.steps, "gpu_launches_per_step": launches_per_step,
            "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu_baseline}
    if args.layers:
        line["invalid"] = f"debug run with {n_layers} layers"
    if kv_bytes == 1:
        line["variant"] = "fp8_e4m3 KV cache: not the BASELINE configuration (bf16 KV), reported next to it"
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
