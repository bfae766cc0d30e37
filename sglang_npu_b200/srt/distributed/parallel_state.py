"""Tensor-parallel group state and the TP collectives the hot path calls
(reference python/sglang/srt/distributed/parallel_state.py:463-567, communication_op.py:11-13).

One process per GPU; the group is `torch.distributed`'s default (NCCL over NVLink/NVSwitch on the
GPU box, gloo in the CPU tests).  At decode sizes the o_proj/down_proj all-reduce is latency bound;
NCCL's in-place ncclAllReduce on the compute stream is CUDA-graph capturable like the reference's
PyNccl path.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist

_TP_GROUP = None
_TP_RANK = 0
_TP_SIZE = 1
_TP_COMM = None  # B200TPComm (symmetric-memory exchange fused with norm + quant) or None -> NCCL all-reduce


def init_tensor_parallel(group=None):
    """Adopt an initialised torch.distributed process group as the TP group."""
    global _TP_GROUP, _TP_RANK, _TP_SIZE
    if not dist.is_initialized():
        _TP_GROUP, _TP_RANK, _TP_SIZE = None, 0, 1
        return
    _TP_GROUP = group if group is not None else dist.group.WORLD
    _TP_RANK = dist.get_rank(_TP_GROUP)
    _TP_SIZE = dist.get_world_size(_TP_GROUP)


def destroy_tensor_parallel():
    global _TP_GROUP, _TP_RANK, _TP_SIZE, _TP_COMM
    if _TP_COMM is not None:
        _TP_COMM.close()
    _TP_GROUP, _TP_RANK, _TP_SIZE, _TP_COMM = None, 0, 1, None


def init_tp_comm(max_tokens: int, hidden: int, dtype, device):
    """Create the fused exchange (csrc/tp_comm.cu) for this TP group; B200_TP_FUSED_EXCHANGE=0 keeps NCCL."""
    global _TP_COMM
    if _TP_SIZE == 1 or os.environ.get("B200_TP_FUSED_EXCHANGE", "1") == "0" or not torch.cuda.is_available():
        return None
    from .b200_comm import B200TPComm

    if _TP_COMM is not None:
        if _TP_COMM.max_tokens >= max_tokens and _TP_COMM.hidden == hidden and _TP_COMM.dtype == dtype:
            return _TP_COMM
        _TP_COMM.close()
    _TP_COMM = B200TPComm(_TP_GROUP, max_tokens, hidden, dtype, device)
    return _TP_COMM


def get_tp_comm():
    return _TP_COMM


def get_tensor_model_parallel_rank() -> int:
    return _TP_RANK


def get_tensor_model_parallel_world_size() -> int:
    return _TP_SIZE


def tensor_model_parallel_all_reduce(x: torch.Tensor) -> torch.Tensor:
    """In-place sum over the TP group (communication_op.py:11-13).  With a symmetric-memory exchange object installed
    (init_tp_comm) a [.., hidden] activation that fits it goes through b200_tp_all_reduce — the counterpart of the
    reference's custom small-message all-reduce (parallel_state.py:463-542 picks it ahead of NCCL the same way);
    everything else (other widths / dtypes, CPU tensors, gloo groups) takes torch.distributed.  The choice depends on
    shapes only, so every rank makes the same one."""
    if _TP_SIZE == 1:
        return x
    comm = _TP_COMM
    if comm is not None and os.environ.get("B200_CUSTOM_ALLREDUCE", "1") != "0" and comm.can_all_reduce(x):
        M = x.numel() // comm.hidden
        comm.partial(M).copy_(x.view(M, comm.hidden))
        x.copy_(comm.all_reduce(M).view_as(x))
        return x
    dist.all_reduce(x, op=dist.ReduceOp.SUM, group=_TP_GROUP)
    return x


def tensor_model_parallel_all_reduce_max(x: torch.Tensor) -> torch.Tensor:
    """In-place max over the TP group (the 8-byte-per-row key exchange of the vocab-parallel greedy sampler)."""
    if _TP_SIZE == 1:
        return x
    dist.all_reduce(x, op=dist.ReduceOp.MAX, group=_TP_GROUP)
    return x


def tensor_model_parallel_all_gather(x: torch.Tensor, dim: int = -1) -> torch.Tensor:
    if _TP_SIZE == 1:
        return x
    if dim < 0:
        dim += x.dim()
    parts = [torch.empty_like(x) for _ in range(_TP_SIZE)]
    dist.all_gather(parts, x.contiguous(), group=_TP_GROUP)
    return torch.cat(parts, dim=dim)
