"""Kernel timeline of ONE CUDA-graph replay of the cfg-2 decode step (torch.profiler / CUPTI):
per-kernel busy time, launch gaps, and a per-kernel-name table.  Diagnostic only — numbers
taken under a profiler are never bench values.

    python tools/step_timeline.py [--layers 32] [--bs 64] [--ctx 4096] [--out gpurun_out/timeline.json]
"""
import argparse
import collections
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sglang_npu_b200 import synth  # noqa: E402
from sglang_npu_b200.srt.layers.quantization.w8a8_int8 import B200W8A8Int8Config  # noqa: E402
from sglang_npu_b200.srt.model_executor.layer_stack import CONFIGS  # noqa: E402
from sglang_npu_b200.srt.model_executor.model_runner import HotPathRunner  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--layers", type=int, default=32)
ap.add_argument("--bs", type=int, default=64)
ap.add_argument("--ctx", type=int, default=4096)
ap.add_argument("--model", default="llama-3.1-8b")
ap.add_argument("--out", default="gpurun_out/timeline.json")
a = ap.parse_args()
dev = torch.device("cuda:0")
cfg = CONFIGS[a.model]
pb = synth.make_paged_batch([a.ctx] * a.bs, 16, seed=1, width=a.ctx + 4, max_reqs=a.bs + 3)
runner = HotPathRunner(cfg, B200W8A8Int8Config(), max_reqs=pb.req_to_token.shape[0], max_total_tokens=pb.num_slots,
                       page_size=16, context_len=a.ctx, device=dev, num_layers=a.layers)
runner.req_to_token_pool.req_to_token.copy_(pb.req_to_token)
for buf in runner.token_to_kv_pool.k_buffer + runner.token_to_kv_pool.v_buffer:
    buf.normal_()
out_loc = torch.stack([pb.req_to_token[pb.req_pool_indices[i], a.ctx - 1] for i in range(a.bs)]).long().to(dev)
ids = torch.randint(0, cfg.vocab_size, (a.bs,), device=dev)
rpi, sl = pb.req_pool_indices.to(dev), pb.seq_lens.to(dev)
runner.capture_decode_graph(a.bs)
for _ in range(3):
    runner.replay_decode(ids, rpi, sl, out_loc)
torch.cuda.synchronize()

from torch.profiler import ProfilerActivity, profile  # noqa: E402

with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    runner.replay_decode(ids, rpi, sl, out_loc)
    torch.cuda.synchronize()

evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and e.time_range is not None]
ks = sorted(((e.time_range.start, e.time_range.end, e.name) for e in evs), key=lambda t: t[0])
if not ks:
    print("no CUDA events captured")
    sys.exit(1)
t0, t1 = ks[0][0], max(k[1] for k in ks)
busy = sum(k[1] - k[0] for k in ks)
agg = collections.defaultdict(lambda: [0, 0.0])
gaps = 0.0
prev_end = None
for s, e, n in ks:
    agg[n[:70]][0] += 1
    agg[n[:70]][1] += e - s
    if prev_end is not None and s > prev_end:
        gaps += s - prev_end
    prev_end = max(prev_end or e, e)
rows = sorted(agg.items(), key=lambda kv: -kv[1][1])
print(f"span {t1 - t0:.1f} us, busy {busy:.1f} us, gaps {gaps:.1f} us, kernels {len(ks)}")
for n, (c, t) in rows[:30]:
    print(f"{n:70s} n={c:4d} avg={t / c:8.2f} us tot={t / 1e3:8.3f} ms {100 * t / (t1 - t0):5.1f}%")
os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
json.dump({"span_us": t1 - t0, "busy_us": busy, "gaps_us": gaps, "n_kernels": len(ks),
           "by_name": [{"name": n, "count": c, "total_us": t} for n, (c, t) in rows]}, open(a.out, "w"), indent=1)
