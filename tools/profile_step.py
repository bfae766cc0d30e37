"""One eager decode step of the cfg-2 workload with few layers — the command profiled under ncu.
    python tools/profile_step.py [--layers 2] [--steps 3] [--bs 64] [--ctx 4096]
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sglang_npu_b200 import kernels, synth  # noqa: E402
from sglang_npu_b200.srt.layers.quantization.w8a8_int8 import B200W8A8Int8Config  # noqa: E402
from sglang_npu_b200.srt.model_executor.layer_stack import CONFIGS  # noqa: E402
from sglang_npu_b200.srt.model_executor.model_runner import HotPathRunner  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--layers", type=int, default=2)
ap.add_argument("--steps", type=int, default=3)
ap.add_argument("--bs", type=int, default=64)
ap.add_argument("--ctx", type=int, default=4096)
ap.add_argument("--model", default="llama-3.1-8b")
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
torch.cuda.synchronize()
torch.cuda.nvtx.range_push("steps")
for _ in range(a.steps):
    logits = runner.forward_decode(ids, rpi, sl, out_loc)
    ids = kernels.argmax(logits)
torch.cuda.synchronize()
torch.cuda.nvtx.range_pop()
print("ok", ids[:4].tolist())
