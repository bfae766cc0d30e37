"""Debug: list mismatches of the fp8 KV write against the golden vectors."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sglang_npu_b200 import kernels
from tests.util import golden_files, load_golden

dev = torch.device("cuda")
for ri, rec in enumerate(load_golden(golden_files("kvwrite_fp8")[0])):
    kc = torch.zeros_like(rec["k_cache_after"], device=dev)
    vc = torch.zeros_like(kc)
    kernels.set_kv_buffer(kc, vc, rec["loc"].to(dev), rec["k"].to(dev), rec["v"].to(dev), rec["k_scale"], rec["v_scale"])
    torch.cuda.synchronize()
    for name, got, exp, src in (("k", kc.cpu(), rec["k_cache_after"], rec["k"]), ("v", vc.cpu(), rec["v_cache_after"], rec["v"])):
        g = got[rec["loc"]].reshape(-1); e = exp[rec["loc"]].reshape(-1); s = src.reshape(-1).float()
        bad = (g != e).nonzero().flatten()
        print(f"rec {ri} {name}: {bad.numel()} / {g.numel()} mismatches; rest-of-cache equal: "
              f"{torch.equal(got, exp) or bad.numel() > 0}")
        for i in bad[:12].tolist():
            print(f"   idx {i} src {s[i].item()!r} exp {e[i].item():#x} got {g[i].item():#x}")
