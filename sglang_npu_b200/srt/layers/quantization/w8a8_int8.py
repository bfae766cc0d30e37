"""`--quantization b200_w8a8_int8` — W8A8-int8 linear method on the B200 kernels.

Drop-in for W8A8Int8Config / W8A8Int8LinearMethod (reference
python/sglang/srt/layers/quantization/w8a8_int8.py:184-394): same parameters (`weight` [N_local,
K_local] int8 ModelWeightParameter, `weight_scale` [N_local, 1] fp32 ChannelQuantScaleParameter), same
`process_weights_after_loading` (weight -> weight.t(), logical [K, N] column-major), same `apply`
contract; the two kernels behind it are csrc/quant.cu (per-token quant) and csrc/gemm_int8.cu
(tcgen05 int8 GEMM; the reference's int8_scaled_mm has no sm_100 build).
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional

import torch
from torch.nn import Parameter

from .... import kernels
from ..parameter import ChannelQuantScaleParameter, ModelWeightParameter
from .base_config import LinearMethodBase, QuantizationConfig, QuantizeMethodBase


class B200W8A8Int8Config(QuantizationConfig):
    """Weight: static, per-channel, symmetric.  Activation: dynamic, per-token, symmetric."""

    def __init__(self, quant_config: Dict[str, Any] = {}):
        super().__init__()
        self.quant_description = quant_config
        self.ignore = list(quant_config.get("ignore", []) or [])

    @classmethod
    def get_supported_act_dtypes(cls) -> List[torch.dtype]:
        return [torch.float16, torch.bfloat16]

    @classmethod
    def get_min_capability(cls) -> int:
        return 100  # tcgen05 kind::i8 (sm_100a)

    @classmethod
    def get_name(cls) -> str:
        return "b200_w8a8_int8"

    @staticmethod
    def get_config_filenames() -> List[str]:
        return []

    @classmethod
    def from_config(cls, config: Dict[str, Any]) -> "B200W8A8Int8Config":
        return cls(config)

    def get_quant_method(self, layer: torch.nn.Module, prefix: str) -> Optional[QuantizeMethodBase]:
        from ..linear import LinearBase

        if isinstance(layer, LinearBase):
            if any(ig in prefix for ig in self.ignore):
                return None
            return B200W8A8Int8LinearMethod(self)
        return None

    def get_scaled_act_names(self) -> List[str]:
        return []


class B200W8A8Int8LinearMethod(LinearMethodBase):
    def __init__(self, quantization_config: B200W8A8Int8Config):
        self.quantization_config = quantization_config

    def create_weights(self, layer: torch.nn.Module, input_size_per_partition: int, output_partition_sizes: List[int],
                       input_size: int, output_size: int, params_dtype: torch.dtype, **extra_weight_attrs):
        weight_loader = extra_weight_attrs.get("weight_loader")
        self.logical_widths = output_partition_sizes
        weight = ModelWeightParameter(
            data=torch.empty(sum(output_partition_sizes), input_size_per_partition, dtype=torch.int8),
            input_dim=1, output_dim=0, weight_loader=weight_loader)
        layer.register_parameter("weight", weight)
        weight_scale = ChannelQuantScaleParameter(
            data=torch.empty((sum(output_partition_sizes), 1), dtype=torch.float32),
            output_dim=0, weight_loader=weight_loader)
        layer.register_parameter("weight_scale", weight_scale)

    def process_weights_after_loading(self, layer: torch.nn.Module) -> None:
        # Idempotent (weight updates re-run it, model_runner.py:731-897): only transpose the [N, K] layout.
        w = layer.weight.data
        if w.stride(-1) == 1 and w.dim() == 2 and not getattr(layer, "_b200_w_transposed", False):
            layer.weight = Parameter(w.t(), requires_grad=False)
            layer._b200_w_transposed = True
        layer.weight_scale = Parameter(layer.weight_scale.data, requires_grad=False)

    def apply(self, layer: torch.nn.Module, x: torch.Tensor, bias: Optional[torch.Tensor] = None,
              out: Optional[torch.Tensor] = None):
        """`out` (optional, beyond the reference signature): write the [M, N] result there — the symmetric buffer of
        the fused TP exchange for row-parallel layers."""
        if isinstance(x, kernels.Int8Activation):  # quantised by the producer kernel (fused rmsnorm / silu-mul)
            return kernels.int8_scaled_mm(x.x_q, layer.weight, x.x_scale, layer.weight_scale, out_dtype=x.out_dtype,
                                          bias=bias, out=out)
        x_q, x_scale = kernels.per_token_quant_int8(x)
        res = kernels.int8_scaled_mm(x_q.view(-1, x_q.shape[-1]), layer.weight, x_scale, layer.weight_scale,
                                     out_dtype=x.dtype, bias=bias, out=out)
        return res.view(x.shape[:-1] + (res.shape[-1],))
