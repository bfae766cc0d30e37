"""Fused TP exchange (csrc/tp_comm.cu) on real GPUs: needs >= 2 devices on the box, one process per GPU
(tools/tp_exchange_check.py under torch.distributed.run).  Skipped on a single-GPU box — there the host-side TP
logic is covered by tests/test_tp_gloo.py (gloo, world_size 2)."""
import json
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_fused_exchange_matches_nccl_path(cuda_device):
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs 2 GPUs (run under gpurun --gpus 2)")
    world = 2
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", str(_free_port()), os.path.join(ROOT, "tools", "tp_exchange_check.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, cwd=ROOT, timeout=900)
    lines = [json.loads(l) for l in res.stdout.splitlines() if l.startswith("{")]
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    checks = [l for l in lines if "ok" in l]
    assert len(checks) >= 9 and all(l["ok"] for l in checks), checks


@pytest.mark.parametrize("tp,M", [(2, 64), (2, 7), (4, 64), (8, 100)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_exchange_kernel_loopback_all_ranks_on_one_gpu(cuda_device, tp, M, dtype):
    """The exchange kernel itself on a single-GPU box: all `tp` ranks run in ONE cooperative launch (blockIdx.y = rank,
    b200_tp_exchange_loopback), each with its own buffer set on the same device, and meet through the same flag protocol
    they use over NVLink (the pointer tables simply hold local addresses).  Checked against the un-fused chain: fp32
    rank-order sum of the partials -> fused_add_rmsnorm -> per-token int8 quant (bit-exact), for the int8 output
    (W8A8-int8 consumers) and the normalised 16-bit output (FP8 / AWQ consumers).  The cooperative launch guarantees
    co-residency of all ranks' CTAs, and the barrier still has its timeout (B200_TP_TIMEOUT_MS): it cannot hang the box."""
    import ctypes

    from sglang_npu_b200 import _cabi, kernels
    from sglang_npu_b200._cabi import dtype_code

    os.environ.setdefault("B200_TP_TIMEOUT_MS", "4000")
    lib = _cabi.load()
    dev = cuda_device
    H, eps, max_blocks = 4096, 1e-5, 16
    g = torch.Generator(device=dev).manual_seed(tp * 100 + M)
    w = (1 + 0.1 * torch.randn(H, generator=g, device=dev)).to(dtype)
    partial = [torch.randn(M, H, generator=g, device=dev).to(dtype) for _ in range(tp)]
    residual0 = torch.randn(M, H, generator=g, device=dev).to(dtype)
    xq = [torch.zeros(M, H, dtype=torch.int8, device=dev) for _ in range(tp)]
    xs = [torch.zeros(M, dtype=torch.float32, device=dev) for _ in range(tp)]
    y = [torch.zeros(M, H, dtype=dtype, device=dev) for _ in range(tp)]
    flags = [torch.zeros(2 * 8 * max_blocks, dtype=torch.int32, device=dev) for _ in range(tp)]
    epoch = [torch.zeros(max_blocks, dtype=torch.int32, device=dev) for _ in range(tp)]
    err = [torch.zeros(64, dtype=torch.int32, device=dev) for _ in range(tp)]
    residual = [residual0.clone() for _ in range(tp)]
    scratch = torch.zeros(tp * 512, dtype=torch.uint8, device=dev)
    arr = lambda ts: (ctypes.c_void_p * tp)(*[t.data_ptr() for t in ts])  # noqa: E731
    stream = torch.cuda.current_stream(dev).cuda_stream
    res_after = None
    for rep in range(2):  # twice: epochs / flags are reused across launches
        rc = lib.b200_tp_exchange_loopback(arr(partial), arr(xq), arr(xs), arr(y), arr(flags), arr(epoch), arr(residual),
                                           w.data_ptr(), M, H, eps, tp, max_blocks, dtype_code(dtype), arr(err),
                                           scratch.data_ptr(), scratch.numel(), stream)
        kernels.check(rc, "b200_tp_exchange_loopback")
        torch.cuda.synchronize(dev)
        for r in range(tp):
            assert int(err[r][0]) == 0, f"rank {r}: barrier timed out (code {int(err[r][0]):#x})"
        # the un-fused chain on the same inputs
        tot = torch.zeros(M, H, dtype=torch.float32, device=dev)
        for p in partial:
            tot += p.float()
        ar = tot.to(dtype)
        per = (M + tp - 1) // tp
        for r in range(tp):
            rows = slice(min(M, r * per), min(M, (r + 1) * per))
            if rows.start == rows.stop:
                continue
            res_ref = (residual0 if rep == 0 else res_after)[rows].clone()
            act, yn = kernels.rmsnorm_quant_int8(ar[rows].clone(), w, eps, residual=res_ref, return_normed=True)
            for dst in range(tp):  # every rank received rank r's rows
                assert torch.equal(xq[dst][rows], act.x_q), f"int8 rows of rank {r} differ on rank {dst}"
                assert torch.equal(xs[dst][rows].view(torch.int32), act.x_scale.view(-1).view(torch.int32))
                assert torch.equal(y[dst][rows].view(torch.int16), yn.view(torch.int16))
            assert torch.equal(residual[r][rows].view(torch.int16), res_ref.view(torch.int16)), "owner's residual rows"
        res_after = torch.cat([residual[r][min(M, r * per): min(M, (r + 1) * per)] for r in range(tp)])


@pytest.mark.parametrize("tp,M", [(2, 64), (4, 7), (8, 100), (8, 3)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_all_reduce_loopback_all_ranks_on_one_gpu(cuda_device, tp, M, dtype):
    """b200_tp_all_reduce's kernel mode (no norm weight: the summed rows go out as they are) through the loopback entry:
    every rank ends up with the fp32 rank-order sum of the partials, rounded once — bit for bit, on every rank, also
    when M < tp leaves ranks without rows, and on a second launch over the same flags."""
    import ctypes

    from sglang_npu_b200 import _cabi, kernels
    from sglang_npu_b200._cabi import dtype_code

    os.environ.setdefault("B200_TP_TIMEOUT_MS", "4000")
    lib = _cabi.load()
    dev = cuda_device
    H, max_blocks = 5120, 16
    g = torch.Generator(device=dev).manual_seed(tp * 1000 + M)
    y = [torch.full((M, H), 7.0, dtype=dtype, device=dev) for _ in range(tp)]
    flags = [torch.zeros(2 * 8 * max_blocks, dtype=torch.int32, device=dev) for _ in range(tp)]
    epoch = [torch.zeros(max_blocks, dtype=torch.int32, device=dev) for _ in range(tp)]
    err = [torch.zeros(64, dtype=torch.int32, device=dev) for _ in range(tp)]
    scratch = torch.zeros(tp * 512, dtype=torch.uint8, device=dev)
    arr = lambda ts: (ctypes.c_void_p * tp)(*[t.data_ptr() for t in ts])  # noqa: E731
    stream = torch.cuda.current_stream(dev).cuda_stream
    for rep in range(2):
        partial = [torch.randn(M, H, generator=g, device=dev).to(dtype) for _ in range(tp)]
        rc = lib.b200_tp_exchange_loopback(arr(partial), None, None, arr(y), arr(flags), arr(epoch), None, None, M, H, 0.0,
                                           tp, max_blocks, dtype_code(dtype), arr(err), scratch.data_ptr(),
                                           scratch.numel(), stream)
        kernels.check(rc, "b200_tp_exchange_loopback (all-reduce mode)")
        torch.cuda.synchronize(dev)
        tot = torch.zeros(M, H, dtype=torch.float32, device=dev)
        for p in partial:
            tot += p.float()
        want = tot.to(dtype)
        for r in range(tp):
            assert int(err[r][0]) == 0, f"rank {r}: barrier timed out (code {int(err[r][0]):#x})"
            assert torch.equal(y[r].view(torch.int16), want.view(torch.int16)), f"rank {r}, launch {rep}"


def test_all_reduce_mode_rejects_fused_arguments(cuda_device):
    """Without a norm weight only the summed rows can be asked for: int8 outputs or a residual are refused (loudly)."""
    import ctypes

    from sglang_npu_b200 import _cabi
    from sglang_npu_b200._cabi import dtype_code

    lib = _cabi.load()
    dev, tp, M, H, mb = cuda_device, 2, 4, 256, 4
    t = lambda *s, dt=torch.bfloat16: [torch.zeros(*s, dtype=dt, device=dev) for _ in range(tp)]  # noqa: E731
    arr = lambda ts: (ctypes.c_void_p * tp)(*[x.data_ptr() for x in ts])  # noqa: E731
    partial, y, res = t(M, H), t(M, H), t(M, H)
    xq, xs = t(M, H, dt=torch.int8), t(M, dt=torch.float32)
    flags, epoch, err = t(2 * 8 * mb, dt=torch.int32), t(mb, dt=torch.int32), t(64, dt=torch.int32)
    scratch = torch.zeros(tp * 512, dtype=torch.uint8, device=dev)
    st = torch.cuda.current_stream(dev).cuda_stream
    rc = lib.b200_tp_exchange_loopback(arr(partial), arr(xq), arr(xs), arr(y), arr(flags), arr(epoch), None, None, M, H, 0.0,
                                       tp, mb, dtype_code(torch.bfloat16), arr(err), scratch.data_ptr(), scratch.numel(), st)
    assert rc != 0
    rc = lib.b200_tp_exchange_loopback(arr(partial), None, None, arr(y), arr(flags), arr(epoch), arr(res), None, M, H, 0.0,
                                       tp, mb, dtype_code(torch.bfloat16), arr(err), scratch.data_ptr(), scratch.numel(), st)
    assert rc != 0
