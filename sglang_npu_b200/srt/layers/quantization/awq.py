"""`--quantization b200_awq` — AWQ-int4 linear method on the B200 fused dequant + tcgen05 GEMM.

Drop-in for AWQConfig / AWQLinearMethod (reference python/sglang/srt/layers/quantization/awq.py:71-146,
:329-418): same config keys (`w_bit`/`bits`, `q_group_size`/`group_size`, `zero_point`,
`modules_to_not_convert`), same parameters (`qweight` [K, N/8] int32 PackedvLLMParameter, `qzeros`
[K/g, N/8] int32, `scales` [K/g, N] GroupQuantScaleParameter) and shape rules (:339-352), same `apply`
contract.  The reference dequantises the whole matrix (`awq_dequantize`) and calls cuBLAS; here
`kernels.awq_gemm` streams the packed words once and dequantises in registers (csrc/awq.cu).  The
reference supports fp16 activations only (:111-112); the kernels also take bf16 (as `awq_dequantize`
does, awq_kernel.cu:209-216).
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional

import torch
from torch.nn import Parameter

from .... import kernels
from ..parameter import GroupQuantScaleParameter, PackedvLLMParameter
from .base_config import LinearMethodBase, QuantizationConfig, QuantizeMethodBase


def is_layer_skipped_awq(prefix: str, modules_to_not_convert: List[str]) -> bool:
    return any(module_name in prefix for module_name in modules_to_not_convert)


class B200AWQConfig(QuantizationConfig):
    def __init__(self, weight_bits: int = 4, group_size: int = 128, zero_point: bool = True,
                 modules_to_not_convert: Optional[List[str]] = None):
        super().__init__()
        self.weight_bits = weight_bits
        self.group_size = group_size
        self.zero_point = zero_point
        self.modules_to_not_convert = modules_to_not_convert or []
        if self.weight_bits != 4:
            raise ValueError(f"Currently, only 4-bit weight quantization is supported for AWQ, but got "
                             f"{self.weight_bits} bits.")
        self.pack_factor = 32 // self.weight_bits

    def __repr__(self) -> str:
        return (f"B200AWQConfig(weight_bits={self.weight_bits}, group_size={self.group_size}, "
                f"zero_point={self.zero_point}, modules_to_not_convert={self.modules_to_not_convert})")

    @classmethod
    def get_name(cls) -> str:
        return "b200_awq"

    @classmethod
    def get_supported_act_dtypes(cls) -> List[torch.dtype]:
        return [torch.half, torch.bfloat16]

    @classmethod
    def get_min_capability(cls) -> int:
        return 100

    @staticmethod
    def get_config_filenames() -> List[str]:
        return ["quant_config.json", "quantize_config.json"]

    @classmethod
    def from_config(cls, config: Dict[str, Any]) -> "B200AWQConfig":
        return cls(cls.get_from_keys(config, ["w_bit", "bits"]),
                   cls.get_from_keys(config, ["q_group_size", "group_size"]),
                   cls.get_from_keys(config, ["zero_point"]),
                   cls.get_from_keys_or(config, ["modules_to_not_convert"], None))

    def get_quant_method(self, layer: torch.nn.Module, prefix: str) -> Optional[QuantizeMethodBase]:
        from ..linear import LinearBase, UnquantizedLinearMethod

        if isinstance(layer, LinearBase):
            if is_layer_skipped_awq(prefix, self.modules_to_not_convert):
                return UnquantizedLinearMethod()
            return B200AWQLinearMethod(self)
        return None

    def get_scaled_act_names(self) -> List[str]:
        return []


class B200AWQLinearMethod(LinearMethodBase):
    def __init__(self, quant_config: B200AWQConfig):
        self.quant_config = quant_config

    def create_weights(self, layer: torch.nn.Module, input_size_per_partition: int, output_partition_sizes: List[int],
                       input_size: int, output_size: int, params_dtype: torch.dtype, **extra_weight_attrs):
        group_size = self.quant_config.group_size if self.quant_config.group_size != -1 else input_size
        if input_size_per_partition % group_size != 0:
            raise ValueError("The input size is not aligned with the quantized weight shape. This can be caused by "
                             "too large tensor parallel size.")
        output_size_per_partition = sum(output_partition_sizes)
        if output_size_per_partition % self.quant_config.pack_factor != 0:
            raise ValueError("The output size is not aligned with the quantized weight shape. This can be caused by "
                             "too large tensor parallel size.")
        weight_loader = extra_weight_attrs.get("weight_loader")
        pf = self.quant_config.pack_factor
        layer.awq_group_size = group_size
        layer.register_parameter("qweight", PackedvLLMParameter(
            data=torch.empty(input_size_per_partition, output_size_per_partition // pf, dtype=torch.int32),
            input_dim=0, output_dim=1, packed_dim=1, packed_factor=pf, weight_loader=weight_loader))
        layer.register_parameter("qzeros", PackedvLLMParameter(
            data=torch.empty(input_size_per_partition // group_size, output_size_per_partition // pf, dtype=torch.int32),
            input_dim=0, output_dim=1, packed_dim=1, packed_factor=pf, weight_loader=weight_loader))
        layer.register_parameter("scales", GroupQuantScaleParameter(
            data=torch.empty(input_size_per_partition // group_size, output_size_per_partition, dtype=params_dtype),
            input_dim=0, output_dim=1, weight_loader=weight_loader))

    def process_weights_after_loading(self, layer: torch.nn.Module) -> None:
        """The reference re-wraps the three tensors (awq.py:396-399).  Here the hook also produces the streaming layout of
        the fused GEMM (kernels.awq_repack: channel-major packed words + {scale, zero} pairs).  The checkpoint-layout
        tensors stay registered — they are what `awq_dequantize` and the parity tooling read — unless
        B200_AWQ_DROP_CHECKPOINT_LAYOUT=1 frees them (a serving process that only ever calls apply())."""
        import os

        layer.qweight = Parameter(layer.qweight.data, requires_grad=False)
        layer.qzeros = Parameter(layer.qzeros.data, requires_grad=False)
        layer.scales = Parameter(layer.scales.data, requires_grad=False)
        if layer.qweight.is_cuda:
            layer.awq_out_features = layer.qweight.shape[-1] * self.quant_config.pack_factor
            qweight_t, sz = kernels.awq_repack(layer.qweight.data, layer.scales.data, layer.qzeros.data)
            layer.register_buffer("qweight_t", qweight_t, persistent=False)
            layer.register_buffer("awq_sz", sz, persistent=False)
            if os.environ.get("B200_AWQ_DROP_CHECKPOINT_LAYOUT", "0") == "1":
                empty = torch.empty(0, dtype=torch.int32, device=qweight_t.device)
                layer.qweight = Parameter(empty, requires_grad=False)
                layer.qzeros = Parameter(empty.clone(), requires_grad=False)
                layer.scales = Parameter(torch.empty(0, dtype=layer.scales.dtype, device=qweight_t.device), requires_grad=False)

    def apply(self, layer: torch.nn.Module, x: torch.Tensor, bias: Optional[torch.Tensor] = None,
              out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """`out` (optional, beyond the reference signature): destination of the [M, N] result (fused TP exchange)."""
        if getattr(layer, "qweight_t", None) is None:
            raise RuntimeError("B200AWQLinearMethod.apply before process_weights_after_loading (or not on a CUDA device)")
        res = kernels.awq_gemm_repacked(x.reshape(-1, x.shape[-1]), layer.qweight_t, layer.awq_sz, layer.awq_group_size,
                                        bias, out=out)
        return res.reshape(x.shape[:-1] + (layer.awq_out_features,))
