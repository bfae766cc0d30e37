"""`--quantization b200_fp8` — block-scaled W8A8-FP8 linear method (and `b200_w8a8_fp8`, per-channel).

Drop-in for Fp8Config / Fp8LinearMethod with `weight_block_size=[128,128]` (reference
python/sglang/srt/layers/quantization/fp8.py:110-187, :228-334 create_weights, :444-487 apply ->
fp8_utils.py:135-210 dispatch) and for W8A8Fp8LinearMethod (w8a8_fp8.py -> fp8_utils.py:510-749).
Parameters are the reference's: `weight` [N, K] e4m3, `weight_scale_inv` [N/128, K/128] fp32
(BlockQuantScaleParameter) or `weight_scale` [N, 1] (per-channel).  The dispatch chain of the
reference (flashinfer / CUTLASS / aiter / DeepGEMM / Triton) collapses to one path:
csrc/quant.cu per-(token,128) quant + csrc/gemm_umma.cu block-scaled tcgen05 GEMM.
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional

import torch
from torch.nn import Parameter

from .... import kernels
from ..parameter import BlockQuantScaleParameter, ChannelQuantScaleParameter, ModelWeightParameter
from .base_config import LinearMethodBase, QuantizationConfig, QuantizeMethodBase


class B200Fp8Config(QuantizationConfig):
    def __init__(self, is_checkpoint_fp8_serialized: bool = True, activation_scheme: str = "dynamic",
                 ignored_layers: Optional[List[str]] = None, weight_block_size: Optional[List[int]] = None):
        super().__init__()
        if not is_checkpoint_fp8_serialized:
            raise NotImplementedError("b200_fp8: only fp8-serialized checkpoints are supported")
        if activation_scheme != "dynamic":
            raise ValueError("b200_fp8: block-wise quantization only supports the dynamic activation scheme")
        weight_block_size = weight_block_size or [128, 128]
        if list(weight_block_size) != [128, 128]:
            raise NotImplementedError(f"b200_fp8: weight_block_size {weight_block_size} (only [128, 128])")
        self.is_checkpoint_fp8_serialized = True
        self.activation_scheme = activation_scheme
        self.ignored_layers = ignored_layers or []
        self.weight_block_size = list(weight_block_size)

    @classmethod
    def get_name(cls) -> str:
        return "b200_fp8"

    @classmethod
    def get_supported_act_dtypes(cls) -> List[torch.dtype]:
        return [torch.bfloat16, torch.half]

    @classmethod
    def get_min_capability(cls) -> int:
        return 100

    @staticmethod
    def get_config_filenames() -> List[str]:
        return []

    @classmethod
    def from_config(cls, config: Dict[str, Any]) -> "B200Fp8Config":
        quant_method = cls.get_from_keys_or(config, ["quant_method"], "fp8")
        return cls(is_checkpoint_fp8_serialized="fp8" in quant_method,
                   activation_scheme=cls.get_from_keys_or(config, ["activation_scheme"], "dynamic"),
                   ignored_layers=cls.get_from_keys_or(config, ["ignored_layers"], None),
                   weight_block_size=cls.get_from_keys_or(config, ["weight_block_size"], None))

    def get_quant_method(self, layer: torch.nn.Module, prefix: str) -> Optional[QuantizeMethodBase]:
        from ..linear import LinearBase

        if isinstance(layer, LinearBase):
            if any(ig in prefix for ig in self.ignored_layers):
                return None
            return B200Fp8LinearMethod(self)
        return None

    def get_scaled_act_names(self) -> List[str]:
        return []


class B200Fp8LinearMethod(LinearMethodBase):
    def __init__(self, quant_config: B200Fp8Config):
        self.quant_config = quant_config
        self.block_quant = True

    def create_weights(self, layer: torch.nn.Module, input_size_per_partition: int, output_partition_sizes: List[int],
                       input_size: int, output_size: int, params_dtype: torch.dtype, **extra_weight_attrs):
        from ...distributed.parallel_state import get_tensor_model_parallel_world_size

        output_size_per_partition = sum(output_partition_sizes)
        weight_loader = extra_weight_attrs.get("weight_loader")
        tp_size = get_tensor_model_parallel_world_size()
        block_n, block_k = self.quant_config.weight_block_size
        # same divisibility rules as fp8.py:246-264
        if tp_size > 1 and input_size // input_size_per_partition == tp_size and input_size_per_partition % block_k != 0:
            raise ValueError(f"Weight input_size_per_partition = {input_size_per_partition} is not divisible by "
                             f"weight quantization block_k = {block_k}.")
        if (tp_size > 1 and output_size // output_size_per_partition == tp_size) or len(output_partition_sizes) > 1:
            for s in output_partition_sizes:
                if s % block_n != 0:
                    raise ValueError(f"Weight output_partition_size = {s} is not divisible by weight quantization "
                                     f"block_n = {block_n}.")
        layer.logical_widths = output_partition_sizes
        layer.input_size_per_partition = input_size_per_partition
        layer.output_size_per_partition = output_size_per_partition
        layer.orig_dtype = params_dtype
        weight = ModelWeightParameter(
            data=torch.empty(output_size_per_partition, input_size_per_partition, dtype=torch.float8_e4m3fn),
            input_dim=1, output_dim=0, weight_loader=weight_loader)
        layer.register_parameter("weight", weight)
        scale = BlockQuantScaleParameter(
            data=torch.empty((output_size_per_partition + block_n - 1) // block_n,
                             (input_size_per_partition + block_k - 1) // block_k, dtype=torch.float32),
            input_dim=1, output_dim=0, weight_loader=weight_loader)
        scale[:] = torch.finfo(torch.float32).min
        layer.register_parameter("weight_scale_inv", scale)
        layer.register_parameter("input_scale", None)

    def process_weights_after_loading(self, layer: torch.nn.Module) -> None:
        # block quant needs no processing on CUDA (fp8.py:336-361); keep plain Parameters
        layer.weight = Parameter(layer.weight.data, requires_grad=False)
        layer.weight_scale_inv = Parameter(layer.weight_scale_inv.data, requires_grad=False)

    def apply(self, layer: torch.nn.Module, x: torch.Tensor, bias: Optional[torch.Tensor] = None,
              out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """`out` (optional, beyond the reference signature): destination of the [M, N] result — the symmetric buffer
        of the fused TP exchange for row-parallel layers."""
        if isinstance(x, kernels.Fp8GroupActivation):  # quantised by the producer kernel (fused rmsnorm / silu-mul)
            return kernels.fp8_blockwise_scaled_mm(x.x_q, layer.weight.t(), x.x_scale, layer.weight_scale_inv.t(),
                                                   out_dtype=x.out_dtype, bias=bias, out=out)
        input_2d = x.view(-1, x.shape[-1])
        q_input, x_scale = kernels.sgl_per_token_group_quant_fp8(input_2d, self.quant_config.weight_block_size[1])
        res = kernels.fp8_blockwise_scaled_mm(q_input, layer.weight.t(), x_scale, layer.weight_scale_inv.t(),
                                              out_dtype=x.dtype, bias=bias, out=out)
        return res.view(*x.shape[:-1], layer.weight.shape[0])


class B200W8A8Fp8Config(QuantizationConfig):
    """Per-channel weights, dynamic per-token activations (reference w8a8_fp8.py)."""

    def __init__(self, is_checkpoint_fp8_serialized: bool = True):
        super().__init__()
        self.is_checkpoint_fp8_serialized = is_checkpoint_fp8_serialized

    @classmethod
    def get_name(cls) -> str:
        return "b200_w8a8_fp8"

    @classmethod
    def get_supported_act_dtypes(cls) -> List[torch.dtype]:
        return [torch.float16, torch.bfloat16]

    @classmethod
    def get_min_capability(cls) -> int:
        return 100

    @staticmethod
    def get_config_filenames() -> List[str]:
        return []

    @classmethod
    def from_config(cls, config: Dict[str, Any]) -> "B200W8A8Fp8Config":
        return cls(True)

    def get_quant_method(self, layer, prefix):
        from ..linear import LinearBase

        return B200W8A8Fp8LinearMethod(self) if isinstance(layer, LinearBase) else None

    def get_scaled_act_names(self) -> List[str]:
        return []


class B200W8A8Fp8LinearMethod(LinearMethodBase):
    def __init__(self, quantization_config: B200W8A8Fp8Config):
        self.quantization_config = quantization_config

    def create_weights(self, layer, input_size_per_partition, output_partition_sizes, input_size, output_size,
                       params_dtype, **extra_weight_attrs):
        weight_loader = extra_weight_attrs.get("weight_loader")
        n = sum(output_partition_sizes)
        layer.register_parameter("weight", ModelWeightParameter(
            data=torch.empty(n, input_size_per_partition, dtype=torch.float8_e4m3fn), input_dim=1, output_dim=0,
            weight_loader=weight_loader))
        layer.register_parameter("weight_scale", ChannelQuantScaleParameter(
            data=torch.empty((n, 1), dtype=torch.float32), output_dim=0, weight_loader=weight_loader))

    def process_weights_after_loading(self, layer) -> None:
        w = layer.weight.data
        if not getattr(layer, "_b200_w_transposed", False):
            layer.weight = Parameter(w.t(), requires_grad=False)  # [K, N] column-major, as w8a8_fp8.py does
            layer._b200_w_transposed = True
        layer.weight_scale = Parameter(layer.weight_scale.data, requires_grad=False)

    def apply(self, layer, x, bias=None, out=None):
        x2 = x.view(-1, x.shape[-1])
        x_q, x_s = kernels.sgl_per_token_quant_fp8(x2)
        res = kernels.fp8_scaled_mm(x_q, layer.weight, x_s, layer.weight_scale, out_dtype=x.dtype, bias=bias, out=out)
        return res.view(*x.shape[:-1], res.shape[-1])
