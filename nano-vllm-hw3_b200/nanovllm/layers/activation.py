"""SiluAndMul with the per-token quantiser fused in.

Mirrors the reference's nanovllm/layers/activation.py:6-14 (``forward(x)``: ``x, y =
x.chunk(2, -1); F.silu(x) * y``).  ``forward_quant`` returns the product already quantised for
``down_proj`` (models/qwen3.py:117-118) as a ``QuantizedActivation``.  One sm_100a kernel
(``ql_silu_mul_quant``), no torch / CPU fallback.
"""
from __future__ import annotations

import torch
from torch import nn

from nanovllm import _ext
from nanovllm.layers.layernorm import QuantizedActivation


class SiluAndMul(nn.Module):

    def __init__(self):
        super().__init__()

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        shape = list(x.shape[:-1]) + [x.shape[-1] // 2]
        _, _, h = _ext.silu_mul_quant(x.reshape(-1, x.shape[-1]), None, True)
        return h.view(shape)

    def forward_quant(self, x: torch.Tensor, dtype: torch.dtype) -> QuantizedActivation:
        q, s, _ = _ext.silu_mul_quant(x.reshape(-1, x.shape[-1]), dtype, False)
        return QuantizedActivation(q, s)
