"""Column / row parallel linears — the callers of `quant_method.apply` and of the TP all-reduce
(reference python/sglang/srt/layers/linear.py:242-415 column, :424-1120 merged/QKV sharding,
:1285-1309 row-parallel forward).  Same constructor arguments, sharding rule and forward contract
(returns `(output, output_bias)`), minus checkpoint-format special cases.
"""
from __future__ import annotations

from typing import List, Optional

import torch
from torch import nn
from torch.nn import Parameter

from ..distributed.parallel_state import (get_tensor_model_parallel_rank, get_tensor_model_parallel_world_size,
                                          tensor_model_parallel_all_gather, tensor_model_parallel_all_reduce)
from .quantization.base_config import LinearMethodBase, QuantizationConfig


def divide(a: int, b: int) -> int:
    assert a % b == 0, f"{a} is not divisible by {b}"
    return a // b


class UnquantizedLinearMethod(LinearMethodBase):
    """bf16/fp16 weights, library GEMM (plain cuBLAS through torch — allowed for un-fused GEMMs)."""

    def create_weights(self, layer, input_size_per_partition, output_partition_sizes, input_size, output_size,
                       params_dtype, **extra):
        # a TP-aware parameter (reference linear.py:181-194 registers input_dim=1 / output_dim=0 on the plain weight),
        # so layers that fall back to this method (quant ignore lists, quant_config=None) shard through the same loaders
        from .parameter import ModelWeightParameter

        weight = ModelWeightParameter(torch.empty(sum(output_partition_sizes), input_size_per_partition, dtype=params_dtype),
                                      weight_loader=extra.get("weight_loader"), input_dim=1, output_dim=0)
        layer.register_parameter("weight", weight)

    def apply(self, layer, x, bias=None, out=None):
        if out is not None and bias is None:
            return torch.matmul(x, layer.weight.t(), out=out.view(*x.shape[:-1], layer.weight.shape[0]))
        res = torch.nn.functional.linear(x, layer.weight, bias)
        return res if out is None else out.copy_(res.view(out.shape))


class LinearBase(nn.Module):
    def __init__(self, input_size: int, output_size: int, skip_bias_add: bool = False,
                 params_dtype: Optional[torch.dtype] = None, quant_config: Optional[QuantizationConfig] = None,
                 prefix: str = ""):
        super().__init__()
        self.input_size = input_size
        self.output_size = output_size
        self.skip_bias_add = skip_bias_add
        self.params_dtype = params_dtype or torch.get_default_dtype()
        self.prefix = prefix
        if quant_config is None:
            self.quant_method = UnquantizedLinearMethod()
        else:
            self.quant_method = quant_config.get_quant_method(self, prefix=prefix) or UnquantizedLinearMethod()


def _block_scale_units(layer, param, offset: int, size: int):
    """Block-quantised scales hold one entry per `block_n` output rows: convert a logical (offset, size) into
    block units (reference linear.py weight_loader_v2: `(x + block_n - 1) // block_n`)."""
    from .parameter import BlockQuantScaleParameter

    if isinstance(param, BlockQuantScaleParameter):
        block_n = layer.quant_method.quant_config.weight_block_size[0]
        return (offset + block_n - 1) // block_n, (size + block_n - 1) // block_n
    return offset, size


class ColumnParallelLinear(LinearBase):
    """Y = XA + b with A split along its output dim: A = [A_1, ..., A_p] (linear.py:242-415)."""

    def __init__(self, input_size: int, output_size: int, bias: bool = True, gather_output: bool = False,
                 skip_bias_add: bool = False, params_dtype=None, quant_config=None,
                 output_sizes: Optional[List[int]] = None, prefix: str = "", tp_rank: Optional[int] = None,
                 tp_size: Optional[int] = None):
        super().__init__(input_size, output_size, skip_bias_add, params_dtype, quant_config, prefix)
        self.gather_output = gather_output
        self.tp_rank = get_tensor_model_parallel_rank() if tp_rank is None else tp_rank
        self.tp_size = get_tensor_model_parallel_world_size() if tp_size is None else tp_size
        self.input_size_per_partition = input_size
        self.output_size_per_partition = divide(output_size, self.tp_size)
        self.output_partition_sizes = [self.output_size_per_partition]
        if hasattr(self, "output_sizes"):
            self.output_partition_sizes = [divide(s, self.tp_size) for s in self.output_sizes]
        self.quant_method.create_weights(self, self.input_size_per_partition, self.output_partition_sizes,
                                         self.input_size, self.output_size, self.params_dtype,
                                         weight_loader=self.weight_loader)
        if bias:
            self.bias = Parameter(torch.zeros(self.output_size_per_partition, dtype=self.params_dtype),
                                  requires_grad=False)
        else:
            self.register_parameter("bias", None)

    def weight_loader(self, param, loaded_weight: torch.Tensor, loaded_shard_id=None):
        if hasattr(param, "load_column_parallel_weight"):
            param.load_column_parallel_weight(loaded_weight, self.tp_rank)
        else:
            shard = param.data.shape[0]
            param.data.copy_(loaded_weight.narrow(0, self.tp_rank * shard, shard))

    def forward(self, input_):
        bias = self.bias if not self.skip_bias_add else None
        output_parallel = self.quant_method.apply(self, input_, bias)
        output = tensor_model_parallel_all_gather(output_parallel) if self.gather_output else output_parallel
        return output, (self.bias if self.skip_bias_add else None)


class MergedColumnParallelLinear(ColumnParallelLinear):
    """Several column-parallel matrices fused along the output dim (gate_up_proj; linear.py:424-…):
    each logical sub-matrix is sharded separately, so rank r holds [gate_r | up_r]."""

    def __init__(self, input_size: int, output_sizes: List[int], bias: bool = True, gather_output: bool = False,
                 skip_bias_add: bool = False, params_dtype=None, quant_config=None, prefix: str = "",
                 tp_rank=None, tp_size=None):
        self.output_sizes = output_sizes
        super().__init__(input_size, sum(output_sizes), bias, gather_output, skip_bias_add, params_dtype,
                         quant_config, prefix=prefix, tp_rank=tp_rank, tp_size=tp_size)

    def weight_loader(self, param, loaded_weight: torch.Tensor, loaded_shard_id: Optional[int] = None):
        assert loaded_shard_id is not None and loaded_shard_id < len(self.output_sizes)
        shard_offset = sum(self.output_sizes[:loaded_shard_id]) // self.tp_size
        shard_size = self.output_sizes[loaded_shard_id] // self.tp_size
        shard_offset, shard_size = _block_scale_units(self, param, shard_offset, shard_size)
        param.load_merged_column_weight(loaded_weight, shard_offset, shard_size, self.tp_rank)


class QKVParallelLinear(ColumnParallelLinear):
    """Fused QKV projection (linear.py:…-1120): q heads are split over TP ranks; KV heads are split when
    tp <= total_kv_heads and replicated (one head per rank) when tp > total_kv_heads."""

    def __init__(self, hidden_size: int, head_size: int, total_num_heads: int,
                 total_num_kv_heads: Optional[int] = None, bias: bool = True, skip_bias_add: bool = False,
                 params_dtype=None, quant_config=None, prefix: str = "", tp_rank=None, tp_size=None):
        self.hidden_size = hidden_size
        self.head_size = head_size
        self.total_num_heads = total_num_heads
        self.total_num_kv_heads = total_num_kv_heads if total_num_kv_heads is not None else total_num_heads
        tp = get_tensor_model_parallel_world_size() if tp_size is None else tp_size
        self.num_heads = divide(self.total_num_heads, tp)
        if tp >= self.total_num_kv_heads:
            self.num_kv_heads = 1
            self.num_kv_head_replicas = divide(tp, self.total_num_kv_heads)
        else:
            self.num_kv_heads = divide(self.total_num_kv_heads, tp)
            self.num_kv_head_replicas = 1
        output_size = (self.num_heads + 2 * self.num_kv_heads) * tp * self.head_size
        self.output_sizes = [self.num_heads * head_size * tp, self.num_kv_heads * head_size * tp,
                             self.num_kv_heads * head_size * tp]
        super().__init__(hidden_size, output_size, bias, False, skip_bias_add, params_dtype, quant_config,
                         prefix=prefix, tp_rank=tp_rank, tp_size=tp_size)

    def weight_loader(self, param, loaded_weight: torch.Tensor, loaded_shard_id: Optional[str] = None):
        assert loaded_shard_id in ("q", "k", "v")
        q_sz = self.num_heads * self.head_size
        kv_sz = self.num_kv_heads * self.head_size
        offset = {"q": 0, "k": q_sz, "v": q_sz + kv_sz}[loaded_shard_id]
        size = q_sz if loaded_shard_id == "q" else kv_sz
        rank = self.tp_rank if loaded_shard_id == "q" else self.tp_rank // self.num_kv_head_replicas
        offset, size = _block_scale_units(self, param, offset, size)
        param.load_merged_column_weight(loaded_weight, offset, size, rank)


class RowParallelLinear(LinearBase):
    """Y = XA + b with A split along its input dim; partial sums are all-reduced (linear.py:1285-1309);
    the bias is added on rank 0 only (:1299)."""

    def __init__(self, input_size: int, output_size: int, bias: bool = True, input_is_parallel: bool = True,
                 skip_bias_add: bool = False, params_dtype=None, reduce_results: bool = True, quant_config=None,
                 prefix: str = "", tp_rank=None, tp_size=None):
        super().__init__(input_size, output_size, skip_bias_add, params_dtype, quant_config, prefix)
        self.input_is_parallel = input_is_parallel
        self.reduce_results = reduce_results
        self.tp_rank = get_tensor_model_parallel_rank() if tp_rank is None else tp_rank
        self.tp_size = get_tensor_model_parallel_world_size() if tp_size is None else tp_size
        self.input_size_per_partition = divide(input_size, self.tp_size)
        self.quant_method.create_weights(self, self.input_size_per_partition, [self.output_size], self.input_size,
                                         self.output_size, self.params_dtype, weight_loader=self.weight_loader)
        if bias:
            self.bias = Parameter(torch.zeros(self.output_size, dtype=self.params_dtype), requires_grad=False)
        else:
            self.register_parameter("bias", None)

    def weight_loader(self, param, loaded_weight: torch.Tensor, loaded_shard_id=None):
        if hasattr(param, "load_row_parallel_weight"):
            param.load_row_parallel_weight(loaded_weight, self.tp_rank)
        else:
            param.data.copy_(loaded_weight)

    def forward(self, input_, can_fuse_mlp_allreduce: bool = False, out=None):
        """can_fuse_mlp_allreduce (reference linear.py:1285): skip the all-reduce, the caller fuses it into the next
        norm.  `out`: destination of the partial sums (the symmetric buffer of B200TPComm)."""
        if self.input_is_parallel:
            input_parallel = input_
        else:
            shard = self.input_size_per_partition
            input_parallel = input_.narrow(-1, self.tp_rank * shard, shard).contiguous()
        bias_ = None if (self.tp_rank > 0 or self.skip_bias_add) else self.bias
        if out is not None:
            output_parallel = self.quant_method.apply(self, input_parallel, bias=bias_, out=out)
        else:
            output_parallel = self.quant_method.apply(self, input_parallel, bias=bias_)
        if self.reduce_results and self.tp_size > 1 and not can_fuse_mlp_allreduce:
            output = tensor_model_parallel_all_reduce(output_parallel)
        else:
            output = output_parallel
        return output, (self.bias if self.skip_bias_add else None)
