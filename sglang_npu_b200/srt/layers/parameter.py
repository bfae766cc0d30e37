"""TP-aware parameter classes used by `create_weights` — the attribute contract of the reference's
python/sglang/srt/layers/parameter.py (input_dim / output_dim / packed_dim / packed_factor and the
column / row / merged-column shard loaders), restated minimally."""
from __future__ import annotations

from typing import Callable, Optional

import torch
from torch.nn import Parameter


class BasevLLMParameter(Parameter):
    def __new__(cls, data: torch.Tensor, **kwargs):
        return super().__new__(cls, data=data, requires_grad=False)

    def __init__(self, data: torch.Tensor, weight_loader: Optional[Callable] = None):
        self._weight_loader = weight_loader

    @property
    def weight_loader(self):
        return self._weight_loader

    def load_full(self, loaded: torch.Tensor):
        assert self.data.shape == loaded.shape, (self.data.shape, loaded.shape)
        self.data.copy_(loaded)


class _ShardedParameter(BasevLLMParameter):
    def __init__(self, data, weight_loader=None, input_dim: Optional[int] = None, output_dim: Optional[int] = None,
                 packed_dim: Optional[int] = None, packed_factor: int = 1, **kw):
        super().__init__(data, weight_loader)
        self._input_dim = input_dim
        self._output_dim = output_dim
        self._packed_dim = packed_dim
        self._packed_factor = packed_factor

    input_dim = property(lambda self: self._input_dim)
    output_dim = property(lambda self: self._output_dim)
    packed_dim = property(lambda self: self._packed_dim)
    packed_factor = property(lambda self: self._packed_factor)

    def _adjust(self, size: int, offset: int, dim: int):
        if self._packed_dim is not None and dim == self._packed_dim:
            return size // self._packed_factor, offset // self._packed_factor
        return size, offset

    def load_column_parallel_weight(self, loaded: torch.Tensor, tp_rank: int):
        """Shard `loaded` (full tensor) along output_dim for this rank."""
        d = self._output_dim
        shard = self.data.shape[d]
        self.data.copy_(loaded.narrow(d, tp_rank * shard, shard))

    def load_row_parallel_weight(self, loaded: torch.Tensor, tp_rank: int):
        d = self._input_dim
        if d is None:  # replicated along K (e.g. per-channel scales of a row-parallel layer)
            self.data.copy_(loaded)
            return
        shard = self.data.shape[d]
        self.data.copy_(loaded.narrow(d, tp_rank * shard, shard))

    def load_merged_column_weight(self, loaded: torch.Tensor, shard_offset: int, shard_size: int, tp_rank: int):
        """`loaded` is the FULL tensor of one logical sub-matrix (q / k / v / gate / up); copy this rank's
        slice into [shard_offset, shard_offset + shard_size) of the fused parameter (logical units)."""
        d = self._output_dim
        size, off = self._adjust(shard_size, shard_offset, d)
        dst = self.data.narrow(d, off, size)
        dst.copy_(loaded.narrow(d, tp_rank * size, size))


class ModelWeightParameter(_ShardedParameter):
    pass


class ChannelQuantScaleParameter(_ShardedParameter):
    pass


class BlockQuantScaleParameter(_ShardedParameter):
    pass


class GroupQuantScaleParameter(_ShardedParameter):
    pass


class PackedvLLMParameter(_ShardedParameter):
    pass
