"""Quantization method registry — the table a maintainer merges into the reference's
BASE_QUANTIZATION_METHODS (python/sglang/srt/layers/quantization/__init__.py:74-119), so that
`--quantization <name>` selects the B200 linear methods."""
from __future__ import annotations

from typing import Dict, Type

from .awq import B200AWQConfig
from .base_config import QuantizationConfig
from .fp8 import B200Fp8Config, B200W8A8Fp8Config
from .w8a8_int8 import B200W8A8Int8Config

B200_QUANTIZATION_METHODS: Dict[str, Type[QuantizationConfig]] = {
    "b200_w8a8_int8": B200W8A8Int8Config,
    "b200_fp8": B200Fp8Config,
    "b200_w8a8_fp8": B200W8A8Fp8Config,
    "b200_awq": B200AWQConfig,
}
# the reference's own names select the same classes when this table replaces its entries
QUANTIZATION_METHODS: Dict[str, Type[QuantizationConfig]] = {
    **B200_QUANTIZATION_METHODS,
    "w8a8_int8": B200W8A8Int8Config,
    "fp8": B200Fp8Config,
    "w8a8_fp8": B200W8A8Fp8Config,
    "awq": B200AWQConfig,
}


def get_quantization_config(quantization: str) -> Type[QuantizationConfig]:
    """quantization/__init__.py:122-131."""
    if quantization not in QUANTIZATION_METHODS:
        raise ValueError(f"Invalid quantization method: {quantization}. Available methods: "
                         f"{list(QUANTIZATION_METHODS.keys())}")
    return QUANTIZATION_METHODS[quantization]
