"""Symmetric-memory exchange for the row-parallel linears: the B200-native replacement for
`tensor_model_parallel_all_reduce` + fused_add_rmsnorm + per_token_quant_int8 on the TP decode / prefill path
(reference python/sglang/srt/layers/linear.py:1303-1304, distributed/communication_op.py:11-13,
distributed/device_communicators/custom_all_reduce.py, layers/flashinfer_comm_fusion.py).

One `B200TPComm` per TP rank.  It owns ONE device allocation, mapped into every peer process with cudaIpc
(handles travel through `torch.distributed.all_gather_object`), laid out as

    partial  [max_tokens, hidden]  act dtype   <- this rank's row-parallel GEMM writes its partial sums here
    xq       [max_tokens, hidden]  int8        <- every rank's exchange kernel writes its token slice here
    xscale   [max_tokens]          fp32
    y        [max_tokens, hidden]  act dtype   <- same, for the un-quantised consumer (final norm -> lm_head)
    flags    [2, 8, max_blocks]    uint32      <- start / end epochs written by the peers
    epoch    [max_blocks]          uint32      <- local launch counter per CTA
    err      [1]                   uint32      <- first barrier timeout of this rank (0 = none), see check()

`reduce_rmsnorm_quant_int8` launches csrc/tp_comm.cu: a two-shot exchange (reduce-scatter by token slice over peer
loads, all-gather of the int8 result over peer stores) with the residual add, RMSNorm and int8 quant in between.
torch is plumbing here (process group for the handle exchange, tensor views over the raw allocation).
"""
from __future__ import annotations

import ctypes
from ctypes import byref, c_void_p
from typing import List, Optional

import torch
import torch.distributed as dist

from ... import _cabi, kernels
from ..._cabi import check, dtype_code, ptr, stream_ptr

MAX_BLOCKS = 512


class _RawCudaBuffer:
    """Minimal __cuda_array_interface__ holder so torch can view a raw device allocation (uint8)."""

    def __init__(self, address: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (address, False), "version": 3,
                                         "strides": None}


def _align(n: int, a: int = 256) -> int:
    return (n + a - 1) // a * a


class B200TPComm:
    def __init__(self, group, max_tokens: int, hidden: int, dtype: torch.dtype, device: torch.device):
        self.group = group
        self.rank = dist.get_rank(group)
        self.tp = dist.get_world_size(group)
        if not 2 <= self.tp <= 8:
            raise RuntimeError(f"B200TPComm: tp size {self.tp} not supported (2..8)")
        self.max_tokens, self.hidden, self.dtype, self.device = max_tokens, hidden, dtype, torch.device(device)
        es = torch.empty((), dtype=dtype).element_size()
        off, o = {}, 0
        for name, nbytes in (("partial", max_tokens * hidden * es), ("xq", max_tokens * hidden),
                             ("xscale", max_tokens * 4), ("y", max_tokens * hidden * es),
                             ("flags", 2 * 8 * MAX_BLOCKS * 4), ("epoch", MAX_BLOCKS * 4), ("err", 256)):
            off[name] = o
            o += _align(nbytes)
        self._off, self.nbytes = off, o
        lib = _cabi.load()
        torch.cuda.set_device(self.device)
        base = c_void_p()
        check(lib.b200_comm_malloc(byref(base), self.nbytes), "b200_comm_malloc")
        self._base = base.value
        handle = (ctypes.c_ubyte * 64)()
        check(lib.b200_comm_ipc_handle(self._base, handle), "b200_comm_ipc_handle")
        handles: List[Optional[bytes]] = [None] * self.tp
        dist.all_gather_object(handles, bytes(handle), group=group)
        self._peer_base: List[int] = []
        for r, h in enumerate(handles):
            if r == self.rank:
                self._peer_base.append(self._base)
                continue
            p = c_void_p()
            buf = (ctypes.c_ubyte * 64).from_buffer_copy(h)
            check(lib.b200_comm_ipc_open(buf, byref(p)), "b200_comm_ipc_open")
            self._peer_base.append(p.value)
        dist.barrier(group=group)
        self._raw = torch.as_tensor(_RawCudaBuffer(self._base, self.nbytes), device=self.device)
        self._ptr_arrays = {name: (c_void_p * self.tp)(*[b + off[name] for b in self._peer_base]) for name in off}

    # ---- local views ------------------------------------------------------------------------------------------------
    def _view(self, name: str, nbytes: int) -> torch.Tensor:
        o = self._off[name]
        return self._raw[o:o + nbytes]

    def partial(self, M: int) -> torch.Tensor:
        """[M, hidden] view the row-parallel GEMM must write its (un-reduced) output into."""
        assert M <= self.max_tokens
        es = self._raw.new_empty(0, dtype=self.dtype).element_size()
        return self._view("partial", M * self.hidden * es).view(self.dtype).view(M, self.hidden)

    def _xq(self, M):
        return self._view("xq", M * self.hidden).view(torch.int8).view(M, self.hidden)

    def _xscale(self, M):
        return self._view("xscale", M * 4).view(torch.float32).view(M, 1)

    def _y(self, M):
        es = self._raw.new_empty(0, dtype=self.dtype).element_size()
        return self._view("y", M * self.hidden * es).view(self.dtype).view(M, self.hidden)

    def rows_of(self, M: int, rank: Optional[int] = None) -> slice:
        """Token slice whose residual rows live on `rank` between exchanges."""
        r = self.rank if rank is None else rank
        per = (M + self.tp - 1) // self.tp
        return slice(min(M, r * per), min(M, (r + 1) * per))

    # ---- the exchange -------------------------------------------------------------------------------------------------
    def reduce_rmsnorm_quant_int8(self, M: int, weight: torch.Tensor, eps: float, residual: Optional[torch.Tensor],
                                  quant: bool = True, normed: bool = False):
        """Consumes `self.partial(M)` of every rank.  Returns an Int8Activation over all M tokens (quant=True) and / or
        the normalised [M, hidden] activation (normed=True).  `residual` ([M, hidden], updated in place) only needs —
        and only keeps — valid rows in `rows_of(M)`."""
        assert weight.dtype == self.dtype and weight.numel() == self.hidden
        assert residual is None or (residual.is_contiguous() and residual.shape == (M, self.hidden))
        lib = _cabi.load()
        pa = self._ptr_arrays
        kernels.check(lib.b200_tp_reduce_rmsnorm_quant_int8(
            pa["partial"], pa["xq"] if quant else None, pa["xscale"] if quant else None, pa["y"] if normed else None,
            pa["flags"], self._base + self._off["epoch"], ptr(residual), ptr(weight), M, self.hidden, float(eps), self.tp,
            self.rank, MAX_BLOCKS, dtype_code(self.dtype), self._base + self._off["err"], stream_ptr(self.device)),
            "b200_tp_reduce_rmsnorm_quant_int8")
        act = kernels.Int8Activation(self._xq(M), self._xscale(M), self.dtype) if quant else None
        if quant and normed:
            return act, self._y(M)
        return act if quant else self._y(M)

    def reduce_rmsnorm(self, M: int, weight: torch.Tensor, eps: float, residual: Optional[torch.Tensor]) -> torch.Tensor:
        """The exchange for linear methods that quantise their own input (FP8, AWQ) or none at all: all-reduce of the
        partial sums + residual add + RMSNorm, every rank receiving the normalised [M, hidden] rows."""
        return self.reduce_rmsnorm_quant_int8(M, weight, eps, residual, quant=False, normed=True)

    def all_reduce(self, M: int) -> torch.Tensor:
        """Plain all-reduce(sum) of every rank's `self.partial(M)`: the [M, hidden] sum (rank-order fp32 sum, one rounding,
        the same bits on every rank) in this rank's `y` buffer — valid until the next exchange through this object.
        The custom small-message all-reduce of the reference (custom_all_reduce.cuh:417-453) without a fused consumer."""
        lib = _cabi.load()
        pa = self._ptr_arrays
        kernels.check(lib.b200_tp_all_reduce(pa["partial"], pa["y"], pa["flags"], self._base + self._off["epoch"], M,
                                             self.hidden, self.tp, self.rank, MAX_BLOCKS, dtype_code(self.dtype),
                                             self._base + self._off["err"], stream_ptr(self.device)),
                      "b200_tp_all_reduce")
        return self._y(M)

    def can_all_reduce(self, x: torch.Tensor) -> bool:
        return (x.is_cuda and x.dtype == self.dtype and x.dim() >= 1 and x.shape[-1] == self.hidden and x.is_contiguous()
                and 0 < x.numel() // self.hidden <= self.max_tokens)

    def check(self) -> None:
        """Synchronise the current stream and raise if an exchange kernel of this rank gave up waiting for a peer
        (B200_TP_TIMEOUT_MS).  Call it where the reference would surface an NCCL timeout: after a step, before results
        are trusted."""
        lib = _cabi.load()
        check(lib.b200_comm_check(self._base + self._off["err"], stream_ptr(self.device)), "b200_comm_check")

    def close(self):
        """Orderly teardown: drain this rank's stream, meet the peers (nobody may still be loading from / storing to
        a mapping that is about to go away), then unmap the peers and free the local allocation."""
        lib = _cabi.load()
        if not self._base:
            return
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)
        torch.cuda.synchronize(self.device)
        for r, b in enumerate(self._peer_base):
            if r != self.rank and b:
                lib.b200_comm_ipc_close(b)
        self._peer_base = []
        dist.barrier(group=self.group)  # every peer has dropped its mapping of this allocation
        lib.b200_comm_free(self._base)
        self._base = 0
