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
