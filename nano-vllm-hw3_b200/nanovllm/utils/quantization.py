"""Quantization ops and module wrappers -- the reference's nanovllm/utils/quantization.py
surface, backed by the sm_100a library instead of Triton / eager torch.

Same names, argument order and assertion behaviour as the reference (the holder classes QLinear /
QTensorLinear / QGroupLinear are RESTATED from :349-401 -- they are the plugin boundary, SURVEY 8b --
and everything they call is new):

    per_row_quant :147-163        per_row_matmul :310-346       QLinear :349-364
    per_group_quant :194-234      per_group_matmul :265-307     QGroupLinear :385-401
    per_tensor_quant :166-191     per_tensor_matmul :237-262    QTensorLinear :367-382
    apply_per_row_quant :463-472  apply_group_quant :487-496    apply_tensor_quant :475-484

Semantics that differ from the reference, on purpose (DESIGN.md "deviations"):
  * bias is added per output column ``y[m,n] += bias[n]``; the reference's Triton epilogue
    adds ``bias[:, None]`` (a row broadcast, :133) -- a latent bug no model run reaches.
  * per_group_matmul / per_tensor_matmul are true W8A8 here (activations quantised per token
    in the same launch); the reference keeps activations in fp32 and re-materialises an fp32
    [N,K] weight every call.  Parity is therefore tolerance-based against the reference
    (its own test thresholds, :583, :615) and exact against oracle/ at the integer level.
  * the ``apply_*`` functions quantise nanovllm.layers.linear modules IN PLACE (TP kept);
    foreign LinearBase objects are wrapped in QLinear & co. exactly as the reference does.
All tensors must live on a B200; there is no CPU or Triton fallback.
"""
from __future__ import annotations

import torch

from nanovllm import _ext
from nanovllm.layers.linear import LinearBase


@torch.no_grad()
def per_row_quant(x: torch.Tensor, dtype: torch.dtype) -> tuple[torch.Tensor, torch.Tensor]:
    assert dtype in [torch.int8, torch.float8_e4m3fn]
    assert x.dim() == 2
    return _ext.quant_rows(x, dtype)


@torch.no_grad()
def per_tensor_quant(x: torch.Tensor, dtype: torch.dtype) -> tuple[torch.Tensor, torch.Tensor]:
    """One scale for the whole tensor; the scale is a 0-dim tensor in x's dtype (bf16)."""
    assert dtype in [torch.int8, torch.float8_e4m3fn]
    return _ext.quant_tensor(x, dtype)


@torch.no_grad()
def per_group_quant(x: torch.Tensor, group_size: int,
                    dtype: torch.dtype) -> tuple[torch.Tensor, torch.Tensor]:
    assert dtype in [torch.int8, torch.float8_e4m3fn]
    assert x.dim() == 2
    assert x.size(1) % group_size == 0, \
        f"K dim ({x.size(1)}) must be divisible by group_size ({group_size})"
    return _ext.quant_groups(x, group_size, dtype)


@torch.no_grad()
def per_row_matmul(x: torch.Tensor, qw: torch.Tensor, sw: torch.Tensor,
                   bias: torch.Tensor | None, dtype: torch.dtype):
    return _ext.linear(x, qw, sw, bias, dtype, 0)


@torch.no_grad()
def per_group_matmul(x: torch.Tensor, qw: torch.Tensor, sw: torch.Tensor,
                     bias: torch.Tensor | None, group_size: int, dtype: torch.dtype):
    return _ext.linear(x, qw, sw, bias, dtype, group_size)


@torch.no_grad()
def per_tensor_matmul(x: torch.Tensor, qw: torch.Tensor, sw: torch.Tensor,
                      bias: torch.Tensor | None, dtype: torch.dtype):
    sw_rows = sw.to(torch.float32).reshape(1).expand(qw.size(0)).contiguous()
    return _ext.linear(x, qw, sw_rows, bias, dtype, 0)


def _bias_of(b):
    return None if b is None else (b.data if isinstance(b, torch.nn.Parameter) else b)


class QLinear(torch.nn.Module):
    """Per-row quantized Linear layer"""

    def __init__(self, dtype, weight: torch.Tensor, bias: torch.Tensor | None):
        super().__init__()
        assert dtype in [torch.float8_e4m3fn, torch.int8]
        self.dtype = dtype
        self.qw, self.sw = per_row_quant(weight, dtype)
        self.bias = bias

    @classmethod
    def from_linear(cls, other: LinearBase, dtype):
        assert other.tp_size == 1, "Not support TP."
        return cls(dtype, other.weight, other.bias)

    def forward(self, x: torch.Tensor):
        return per_row_matmul(x, self.qw, self.sw, _bias_of(self.bias), self.dtype)


class QTensorLinear(torch.nn.Module):
    """Per-tensor quantized Linear layer"""

    def __init__(self, dtype, weight: torch.Tensor, bias: torch.Tensor | None):
        super().__init__()
        assert dtype in [torch.float8_e4m3fn, torch.int8]
        self.dtype = dtype
        self.qw, self.sw = per_tensor_quant(weight, dtype)
        self._sw_rows = self.sw.to(torch.float32).reshape(1).expand(self.qw.size(0)).contiguous()
        self.bias = bias

    @classmethod
    def from_linear(cls, other: LinearBase, dtype):
        assert other.tp_size == 1, "Not support TP."
        return cls(dtype, other.weight, other.bias)

    def forward(self, x: torch.Tensor):
        return _ext.linear(x, self.qw, self._sw_rows, _bias_of(self.bias), self.dtype, 0)


class QGroupLinear(torch.nn.Module):
    """Per-group quantized Linear layer"""

    def __init__(self, dtype, weight: torch.Tensor, bias: torch.Tensor | None, group_size: int):
        super().__init__()
        assert dtype in [torch.float8_e4m3fn, torch.int8]
        self.dtype = dtype
        self.group_size = group_size
        self.qw, self.sw = per_group_quant(weight, group_size, dtype)
        self.bias = bias

    @classmethod
    def from_linear(cls, other: LinearBase, dtype, group_size: int):
        assert other.tp_size == 1, "Not support TP."
        return cls(dtype, other.weight, other.bias, group_size)

    def forward(self, x: torch.Tensor):
        return per_group_matmul(x, self.qw, self.sw, _bias_of(self.bias), self.group_size, self.dtype)


# ------------------------------------------------------------------ model-level appliers

_PROJS = (("self_attn", "qkv_proj"), ("self_attn", "o_proj"), ("mlp", "gate_up_proj"),
          ("mlp", "down_proj"))


def _apply(model, quant_type: str, dtype, group_size: int, wrap):
    for layer in model.model.layers:
        for parent_name, name in _PROJS:
            parent = getattr(layer, parent_name)
            lin = getattr(parent, name)
            if hasattr(lin, "quantize_"):          # our TP classes: quantise in place, TP kept
                lin.quantize_(dtype, quant_type, group_size)
            else:                                  # a foreign LinearBase: wrap like the reference
                setattr(parent, name, wrap(lin))
    if torch.cuda.is_available():
        torch.cuda.empty_cache()


def apply_per_row_quant(model, dtype):
    """Apply per-row quantization to model"""
    _apply(model, "per_row", dtype, 0, lambda lin: QLinear.from_linear(lin, dtype))


def apply_tensor_quant(model, dtype):
    """Apply per-tensor quantization to model"""
    _apply(model, "per_tensor", dtype, 0, lambda lin: QTensorLinear.from_linear(lin, dtype))


def apply_group_quant(model, dtype, group_size: int):
    """Apply per-group quantization to model"""
    _apply(model, "per_group", dtype, group_size,
           lambda lin: QGroupLinear.from_linear(lin, dtype, group_size))


def apply_quantization(model, linear_dtype, quant_type: str | None = None, group_size: int = 128):
    """The dispatch of ModelRunner.__init__ (engine/model_runner.py:36-57) as one function, for the
    branches that are in scope: "per_tensor", "per_row", "per_group", and ``quant_type=None`` with a
    non-bf16 ``linear_dtype`` (per-row, the reference's default rule).

    Two branches differ from the reference ON PURPOSE: ``quant_type="smoothquant"`` goes to torchao
    there (:43-46; third-party arithmetic no reference test pins, out of scope) and any
    other unknown string is silently ignored there (:47-54 fall through).  Both raise ValueError
    here: running a bf16 model when a quantised one was asked for is not a result worth hiding."""
    if quant_type is not None:
        if quant_type == "per_tensor":
            apply_tensor_quant(model, linear_dtype)
        elif quant_type == "per_row":
            apply_per_row_quant(model, linear_dtype)
        elif quant_type == "per_group":
            apply_group_quant(model, linear_dtype, group_size)
        else:
            raise ValueError(f"Unknown quant_type: {quant_type}")
    elif linear_dtype != torch.bfloat16:
        apply_per_row_quant(model, linear_dtype)
