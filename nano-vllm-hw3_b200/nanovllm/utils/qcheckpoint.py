"""Serialised quantized checkpoints.

The reference never stores quantized weights: every start-up loads bf16 safetensors
(nanovllm/utils/loader.py:12-28) and quantises on the GPU (utils/quantization.py:463-496).  This
module writes the quantized state of the TP linear layers once and loads it straight into
(``qw``, ``sw``) -- no bf16 copy, no on-GPU quantiser at start-up -- sharding the scales together
with the codes:

    <name>.qweight  uint8 view of the int8 / e4m3 codes, [N, K]   (TP=1 layout)
    <name>.scale    fp32 [N] (per_row) or [N, K/g] (per_group)
    <name>.bias     bf16 [N] (optional)
    metadata        {"<name>": "int8|fp8_e4m3,per_row|per_group,<group_size>"}

Sharding on load (``tp_dim`` as in the reference's layer classes, linear.py:54-153):
    column-parallel (tp_dim 0): codes and scales narrow along dim 0 (row-local scales commute
        with the split; merged QKV / gate-up layers shard each sub-matrix separately);
    row-parallel (tp_dim 1): codes narrow along K; per-row scales [N] are replicated (a scale
        computed over the full row is valid for every K slice: y = sum_r acc_r * sx_r * sw),
        per-group scales narrow along dim 1 (needs (K/tp) % g == 0).
"""
from __future__ import annotations

import json

import torch
from safetensors import safe_open
from safetensors.torch import save_file

_DT = {"int8": torch.int8, "fp8_e4m3": torch.float8_e4m3fn}
_DT_NAME = {v: k for k, v in _DT.items()}


def save_quantized(named_layers: dict, path: str) -> None:
    """``named_layers``: {name: quantized LinearBase at TP=1}.  Writes one safetensors file."""
    tensors, meta = {}, {}
    for name, layer in named_layers.items():
        assert layer.is_quantized and layer.tp_size == 1, "checkpoints are written from a TP=1 model"
        assert layer.quant_type in ("per_row", "per_group")
        tensors[f"{name}.qweight"] = layer.qw.view(torch.uint8).cpu().contiguous()
        tensors[f"{name}.scale"] = layer.sw.float().cpu().contiguous()
        if layer.bias is not None:
            tensors[f"{name}.bias"] = layer.bias.data.cpu().contiguous()
        meta[name] = f"{_DT_NAME[layer.qdtype]},{layer.quant_type},{layer.group_size}"
    save_file(tensors, path, metadata={"qlinear_b200": json.dumps(meta)})


def _sub_ranges(layer, n_full: int):
    """Row ranges [(start, size)] of the full [N, K] matrix that this rank's column-parallel shard
    is made of: one range for a plain layer, one per packed sub-matrix for merged layers."""
    tp, rank = layer.tp_size, layer.tp_rank
    if hasattr(layer, "output_sizes"):                          # MergedColumnParallelLinear
        sizes = list(layer.output_sizes)
    elif hasattr(layer, "num_kv_heads"):                        # QKVParallelLinear
        sizes = [layer.num_heads * tp * layer.head_size, layer.num_kv_heads * tp * layer.head_size,
                 layer.num_kv_heads * tp * layer.head_size]
    else:
        sizes = [n_full]
    assert sum(sizes) == n_full
    out, off = [], 0
    for s in sizes:
        out.append((off + rank * (s // tp), s // tp))
        off += s
    return out


def load_quantized(named_layers: dict, path: str, device="cuda") -> None:
    """Fill (``qw``, ``sw``, bias) of every layer in ``named_layers`` (any TP size) from ``path``."""
    with safe_open(path, "pt", "cpu") as f:
        meta = json.loads(f.metadata()["qlinear_b200"])
        for name, layer in named_layers.items():
            dt_name, quant_type, g = meta[name].split(",")
            dtype, g = _DT[dt_name], int(g)
            qw, sw = f.get_tensor(f"{name}.qweight"), f.get_tensor(f"{name}.scale")
            bias = f.get_tensor(f"{name}.bias") if f"{name}.bias" in f.keys() else None
            # the file must describe THIS layer (full, TP=1 shapes): a mismatched or partial file
            # must not load silently
            tp = layer.tp_size
            ref = layer.weight if getattr(layer, "weight", None) is not None else layer.qw
            n_full = ref.shape[0] * (tp if layer.tp_dim == 0 else 1)
            k_full = ref.shape[1] * (tp if layer.tp_dim == 1 else 1)
            if tuple(qw.shape) != (n_full, k_full):
                raise ValueError(f"{name}: qweight {tuple(qw.shape)} does not match the layer ({n_full}, {k_full})")
            n_sw = n_full * (k_full // g) if quant_type == "per_group" else n_full
            if sw.dtype != torch.float32 or sw.numel() != n_sw or sw.shape[0] != n_full:
                raise ValueError(f"{name}: scale {tuple(sw.shape)} / {sw.dtype} does not match "
                                 f"{n_sw} fp32 values in {n_full} rows")
            if layer.bias is not None and bias is None:
                raise ValueError(f"{name}: the layer has a bias but the checkpoint has none")
            if bias is not None and bias.numel() != n_full:
                raise ValueError(f"{name}: bias has {bias.numel()} entries, the layer {n_full}")
            if layer.tp_size > 1 and layer.tp_dim == 0:
                rows = torch.cat([torch.arange(a, a + n) for a, n in _sub_ranges(layer, qw.shape[0])])
                qw, sw = qw[rows], sw[rows]
                bias = bias[rows] if bias is not None else None     # column-parallel bias shards with N
            elif layer.tp_size > 1 and layer.tp_dim == 1:
                k = qw.shape[1] // layer.tp_size
                qw = qw[:, layer.tp_rank * k:(layer.tp_rank + 1) * k]
                if quant_type == "per_group":
                    assert k % g == 0, f"{name}: K shard {k} not a multiple of group_size {g}"
                    sw = sw[:, layer.tp_rank * (k // g):(layer.tp_rank + 1) * (k // g)]
            layer.qw = qw.contiguous().to(device).view(dtype)
            layer.sw = sw.contiguous().to(device)
            layer._sw_kernel = layer.sw
            layer.qdtype, layer.quant_type = dtype, quant_type
            layer.group_size = g if quant_type == "per_group" else 0
            if layer.bias is not None and bias is not None:
                layer.bias.data.copy_(bias)                          # row-parallel: replicated (rank 0 adds it)
            layer.weight = None
