"""Generate golden vectors for the quantised linears by running the REFERENCE'S OWN known-answer functions.

    python tests/golden/make_golden_quant.py          (this container only: reads /root/reference)

The reference's quant package cannot be imported here (transformers-5.5 dataclass error, absent
sgl_kernel wheel), but the functions its tests use as THEIR oracle are plain torch.  They are lifted out of the
reference sources with `ast` (function definitions only — no module import, no reference code is copied into
this repo) and exec'd on seeded inputs; inputs + outputs are stored as small fixtures:

  quant_fp8_group.pt    native_per_token_group_quant_fp8       python/sglang/test/test_block_fp8.py:24-49
  quant_fp8_block_mm.pt native_w8a8_block_fp8_matmul           python/sglang/test/test_block_fp8.py:274-330
                        baseline_scaled_mm (group_broadcast)   sgl-kernel/tests/test_fp8_blockwise_gemm.py:21-61
  quant_fp8_chan.pt     torch_scaled_mm                        sgl-kernel/tests/test_fp8_gemm.py:6-14
                        torch_per_token_quant_fp8              sgl-kernel/tests/test_per_token_quant_fp8.py:14-22
  quant_awq.pt          awq_dequantize_torch                   sgl-kernel/tests/test_awq_dequant.py:9-61
  quant_int8.pt         per_token_quant_int8, native_w8a8_per_token_matmul   test/srt/cpu/utils.py:24-32, :91-115
                        torch_scaled_mm                        sgl-kernel/tests/test_int8_gemm.py:13-19

tests/test_quant_golden_cpu.py holds oracle/quant_oracle.py to these fixtures on the CPU; the GPU tests compare the
CUDA kernels with the same fixtures.  The fixtures travel to the GPU box; the reference does not.
"""
import ast
import os
import sys
from typing import Optional, Tuple, Type  # noqa: F401  (names the lifted functions' annotations use)

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


def lift(rel_path: str, *names: str) -> dict:
    """exec the named top-level function definitions of a reference source file in a torch-only namespace."""
    path = os.path.join(REF, rel_path)
    tree = ast.parse(open(path).read(), filename=path)
    keep = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in names]
    missing = set(names) - {n.name for n in keep}
    assert not missing, f"{rel_path}: {missing} not found"
    ns = {"torch": torch, "Optional": Optional, "Tuple": Tuple, "Type": Type}
    exec(compile(ast.Module(body=keep, type_ignores=[]), path, "exec"), ns)
    return {n: ns[n] for n in names}


def e4m3(t: torch.Tensor) -> torch.Tensor:
    return t.clamp(-448.0, 448.0).to(torch.float8_e4m3fn)


def gen_fp8_group():
    f = lift("python/sglang/test/test_block_fp8.py", "native_per_token_group_quant_fp8")["native_per_token_group_quant_fp8"]
    cases = []
    g = torch.Generator().manual_seed(1234)
    for dtype in (torch.bfloat16, torch.float16, torch.float32):
        for (m, k) in ((7, 512), (33, 1024)):
            x = (torch.randn(m, k, generator=g) * 3).to(dtype)
            x[0, :128] = 0  # all-zero group: the eps clamp decides the scale
            x[1, 128:256] *= 1e-6  # tiny group
            x[2, 0] = 1000.0  # dominated group
            q, s = f(x.contiguous(), 128)
            cases.append(dict(x=x, group=128, q=q.view(torch.uint8), s=s, dtype=str(dtype)))
    return cases


def gen_fp8_block_mm():
    native = lift("python/sglang/test/test_block_fp8.py", "native_w8a8_block_fp8_matmul")["native_w8a8_block_fp8_matmul"]
    base = lift("sgl-kernel/tests/test_fp8_blockwise_gemm.py", "baseline_scaled_mm")["baseline_scaled_mm"]
    cases = []
    g = torch.Generator().manual_seed(4321)
    for (m, n, k) in ((1, 128, 512), (5, 256, 384), (83, 512, 1024), (130, 384, 2048)):
        a = e4m3((torch.rand(m, k, generator=g) - 0.5) * 2 * 448)
        b = e4m3((torch.rand(n, k, generator=g) - 0.5) * 2 * 448)
        # both sign conventions the reference tests use: randn*1e-3 (signed) and rand (positive)
        sa = torch.randn(m, k // 128, generator=g) * 1e-3
        sb = torch.randn(n // 128, k // 128, generator=g) * 1e-3
        for out_dtype in (torch.bfloat16, torch.float16, torch.float32):
            o_native = native(a, b, sa, sb, [128, 128], output_dtype=out_dtype)
            # baseline_scaled_mm takes b [K, N] and scale_b [K/128, N/128]
            o_base = base(a, b.t(), sa, sb.t(), out_dtype)
            cases.append(dict(a=a.view(torch.uint8), b=b.view(torch.uint8), sa=sa, sb=sb, out_native=o_native,
                              out_baseline=o_base, out_dtype=str(out_dtype)))
    return cases


def gen_fp8_chan():
    mm = lift("sgl-kernel/tests/test_fp8_gemm.py", "torch_scaled_mm")["torch_scaled_mm"]
    quant = lift("sgl-kernel/tests/test_per_token_quant_fp8.py", "torch_per_token_quant_fp8")["torch_per_token_quant_fp8"]
    cases = []
    g = torch.Generator().manual_seed(99)
    for (m, n, k, with_bias) in ((1, 128, 512, False), (17, 256, 1024, True), (128, 512, 512, True)):
        a = e4m3((torch.rand(m, k, generator=g) - 0.5) * 2 * 448)
        b = e4m3((torch.rand(n, k, generator=g) - 0.5) * 2 * 448)
        sa = torch.randn(m, generator=g) * 1e-3
        sb = torch.randn(n, generator=g) * 1e-3
        for out_dtype in (torch.bfloat16, torch.float16):
            bias = torch.randn(n, generator=g).to(out_dtype) if with_bias else None
            o = mm(a, b.t(), sa, sb, out_dtype, bias)
            cases.append(dict(kind="mm", a=a.view(torch.uint8), b=b.view(torch.uint8), sa=sa, sb=sb, bias=bias, out=o,
                              out_dtype=str(out_dtype)))
    for dtype in (torch.bfloat16, torch.float16):
        x = (torch.randn(19, 1368, generator=g) * 2).to(dtype)
        x[3] *= 1e-4
        # the kernel's scale: absmax / 448 per token (per_token_quant_fp8.cu); the test function takes it as `inv_scale`
        s = x.float().abs().amax(dim=-1) / 448.0
        q = quant(x, s)
        cases.append(dict(kind="quant", x=x, s=s, q=q.view(torch.uint8), dtype=str(dtype)))
    return cases


def gen_awq():
    f = lift("sgl-kernel/tests/test_awq_dequant.py", "reverse_awq_order", "awq_dequantize_torch")
    # awq_dequantize_torch calls reverse_awq_order through its globals: both live in one namespace above
    deq = f["awq_dequantize_torch"]
    cases = []
    g = torch.Generator().manual_seed(7)
    for (rows, cols8, group, dtype) in ((128, 16, 128, torch.float16), (256, 72, 128, torch.float16),
                                        (512, 32, -1, torch.bfloat16), (384, 64, 128, torch.bfloat16),
                                        (256, 16, 64, torch.float16)):
        gs = rows if group == -1 else group
        qw = torch.randint(0, torch.iinfo(torch.int32).max, (rows, cols8), dtype=torch.int32, generator=g)
        qz = torch.randint(0, torch.iinfo(torch.int32).max, (rows // gs, cols8), dtype=torch.int32, generator=g)
        # top nibble set too (the reference's randint never sets bit 31): flip the sign bit of a few words
        qw[::3, ::2] |= torch.tensor(-(1 << 31), dtype=torch.int32)
        qz[:, 1::2] |= torch.tensor(-(1 << 31), dtype=torch.int32)
        sc = torch.rand(rows // gs, cols8 * 8, generator=g).to(dtype)
        w = deq(qw, sc, qz, group)
        cases.append(dict(qweight=qw, scales=sc, qzeros=qz, group=gs, w=w))
    return cases


def gen_int8():
    u = lift("test/srt/cpu/utils.py", "per_token_quant_int8", "native_w8a8_per_token_matmul")
    mm = lift("sgl-kernel/tests/test_int8_gemm.py", "torch_scaled_mm")["torch_scaled_mm"]
    cases = []
    g = torch.Generator().manual_seed(1234)
    for (m, n, k, with_bias) in ((1, 128, 512, False), (8, 512, 544, True), (65, 256, 1024, True)):
        x = (torch.randn(m, k, generator=g) / 10).to(torch.bfloat16)
        x[0, 0] = 0.5 * x[0].abs().max()  # exact .5 tie candidates for the rounding rule
        w = torch.randint(-127, 128, (n, k), generator=g, dtype=torch.int32).to(torch.int8)
        ws = torch.rand(n, generator=g) * 1e-2
        bias = torch.randn(n, generator=g).to(torch.bfloat16) if with_bias else None
        xq, xs = u["per_token_quant_int8"](x)
        out = u["native_w8a8_per_token_matmul"](xq, w, xs, ws, bias, torch.bfloat16)
        out2 = mm(xq, w.t(), xs.view(-1), ws, torch.bfloat16, bias)
        cases.append(dict(x=x, w=w, ws=ws, bias=bias, xq=xq, xs=xs, out_native=out, out_scaled_mm=out2))
    return cases


def main():
    if not os.path.isdir(REF):
        sys.exit("needs /root/reference (run in the build container)")
    torch.set_num_threads(1)  # fixed fp32 summation order for the matmuls
    for name, fn in (("quant_fp8_group", gen_fp8_group), ("quant_fp8_block_mm", gen_fp8_block_mm),
                     ("quant_fp8_chan", gen_fp8_chan), ("quant_awq", gen_awq), ("quant_int8", gen_int8)):
        cases = fn()
        path = os.path.join(HERE, name + ".pt")
        torch.save(cases, path)
        print(f"{name}: {len(cases)} cases, {os.path.getsize(path) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
