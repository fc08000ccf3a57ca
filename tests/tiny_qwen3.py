"""Test-only miniature of the reference's Qwen3 (nanovllm/models/qwen3.py) -- NOT product code.

It exists so that tests can execute the plugin boundary the way the reference engine does:
``apply_quantization(model, linear_dtype, quant_type, group_size)`` (model_runner.py:36-57)
walks ``model.model.layers[i].self_attn.qkv_proj / .o_proj`` and ``.mlp.gate_up_proj /
.down_proj`` (utils/quantization.py:463-496), so the attribute layout below is the reference's
(qwen3.py:46-57, 102-111, 126-146, 167-178, 201-206).  Qwen3-0.6B widths, ``n_layers`` decoder
layers, seeded random-init weights, an untied random ``lm_head``.

What is NOT the product path and only lives here: the embedding lookup, RoPE
(rotary_embedding.py:6-14), the head-wise q/k norms' reshape, and attention, which is a plain
torch causal SDPA (the reference uses flash-attn, attention.py:78-96 -- out of scope).
RMSNorm / SiluAndMul are the package's mirrors of the reference layers, and the four
projections are whatever ``linear_factory`` builds:

  * ``tp_linears``       the package's QKVParallelLinear / RowParallelLinear / MergedColumn...
                         (quantised in place by ``apply_*``, tensor parallel kept);
  * ``foreign_linears``  a stand-in for a *foreign* ``LinearBase`` (weight, bias, tp_size = 1,
                         no ``quantize_``), which ``apply_*`` must wrap in QLinear / QGroupLinear /
                         QTensorLinear exactly as the reference does (:358-361, :376-379, :394-398);
  * ``OracleLinear``     the CPU oracle (oracle/) as the linear: the restated reference chain.

All models built with the same seed hold identical weights; the non-linear ops are the same
GPU code in every variant, so any difference in logits comes from the linears alone.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F
from torch import nn

from oracle import oracle as orc

CFG_06B = dict(hidden=1024, inter=3072, heads=16, kv=8, hd=128, eps=1e-6, rope_theta=1e6)


# ------------------------------------------------------------------ the four projections

class ForeignLinear(nn.Module):
    """What the reference's own LinearBase looks like to ``apply_*`` (linear.py:12-34): a bf16
    ``weight [N,K]``, ``bias``, ``tp_size`` -- and no ``quantize_`` method."""

    def __init__(self, in_features: int, out_features: int):
        super().__init__()
        self.tp_dim, self.tp_rank, self.tp_size = None, 0, 1
        self.weight = nn.Parameter(torch.empty(out_features, in_features))
        self.bias = None

    def forward(self, x):
        return F.linear(x, self.weight, self.bias)


def tp_linears(cfg):
    from nanovllm.layers.linear import (MergedColumnParallelLinear, QKVParallelLinear,
                                        RowParallelLinear)
    return dict(
        qkv=lambda: QKVParallelLinear(cfg["hidden"], cfg["hd"], cfg["heads"], cfg["kv"]),
        o=lambda: RowParallelLinear(cfg["heads"] * cfg["hd"], cfg["hidden"]),
        gate_up=lambda: MergedColumnParallelLinear(cfg["hidden"], [cfg["inter"]] * 2),
        down=lambda: RowParallelLinear(cfg["inter"], cfg["hidden"]))


def foreign_linears(cfg):
    qkv_n = (cfg["heads"] + 2 * cfg["kv"]) * cfg["hd"]
    return dict(
        qkv=lambda: ForeignLinear(cfg["hidden"], qkv_n),
        o=lambda: ForeignLinear(cfg["heads"] * cfg["hd"], cfg["hidden"]),
        gate_up=lambda: ForeignLinear(cfg["hidden"], 2 * cfg["inter"]),
        down=lambda: ForeignLinear(cfg["inter"], cfg["hidden"]))


class OracleLinear(nn.Module):
    """A linear whose forward is the CPU oracle.  ``semantics``:

    * ``"w8a8"``      what the engine computes, restated: per-token activation quantiser
                      (_per_row_quant_kernel, quantization.py:13-51) + the scaled GEMM
                      (_per_row_post_process_mm :54-144 for per-row / per-tensor scales,
                      the per-128-K-block promotion for per-group);
    * ``"reference"`` what the REFERENCE computes for this quant_type: per_row -> the same
                      Triton chain; per_group / per_tensor -> weight-only dequant + fp32
                      matmul on the unquantised activations (per_group_matmul :265-307,
                      per_tensor_matmul :237-262).
    """

    def __init__(self, weight_bf16: torch.Tensor, dtype, quant_type: str, group_size: int,
                 semantics: str):
        super().__init__()
        assert semantics in ("w8a8", "reference")
        self.fmt = orc.fmt_of(dtype)
        self.quant_type, self.semantics = quant_type, semantics
        w = orc.t2n(weight_bf16)
        N, K = w.shape
        if quant_type == "per_row":
            self.qw, self.sw = orc.quant_rows(w, self.fmt)
            self.g = 0
        elif quant_type == "per_group":
            self.qw, self.sw = orc.quant_groups(w, group_size, self.fmt)
            self.g = group_size
        elif quant_type == "per_tensor":
            self.qw, sc = orc.quant_tensor(w, self.fmt)
            self.sw = np.full((N,), sc, np.float32)
            self.g = 0
        else:
            raise ValueError(quant_type)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        shape = list(x.shape[:-1]) + [self.qw.shape[0]]
        xb = orc.t2n(x.reshape(-1, x.shape[-1]))
        K = xb.shape[1]
        if self.semantics == "reference" and self.quant_type != "per_row":
            g = self.g or K
            sw = self.sw if self.g else self.sw.reshape(-1, 1)
            y = orc.dequant_matmul_groups(xb, self.qw, np.ascontiguousarray(sw), None, g, self.fmt)
        else:
            qx, sx = orc.quant_rows(xb, self.fmt)
            if self.g:
                y = orc.gemm_groups(qx, sx, self.qw, self.sw, None, self.g, self.fmt)
            else:
                y = orc.gemm_rows(qx, sx, self.qw, self.sw, None, self.fmt)
        return orc.n2t(y, torch.bfloat16).to(x.device).view(shape)


# ------------------------------------------------------------------ the model

class TinyAttention(nn.Module):
    def __init__(self, cfg, lin, norm_cls):
        super().__init__()
        self.h, self.kv, self.hd = cfg["heads"], cfg["kv"], cfg["hd"]
        self.qkv_proj = lin["qkv"]()
        self.o_proj = lin["o"]()
        self.q_norm = norm_cls(self.hd, eps=cfg["eps"])
        self.k_norm = norm_cls(self.hd, eps=cfg["eps"])
        self.theta = cfg["rope_theta"]

    def _rope(self, positions, x):
        # rotary_embedding.py:6-14,22-33: neox-style halves, fp32 math, cast back
        half = self.hd // 2
        inv = 1.0 / (self.theta ** (torch.arange(0, self.hd, 2, dtype=torch.float, device=x.device) / self.hd))
        f = positions.float()[:, None] * inv[None, :]
        cos, sin = f.cos()[:, None, :], f.sin()[:, None, :]
        x1, x2 = x.float()[..., :half], x.float()[..., half:]
        return torch.cat((x1 * cos - x2 * sin, x2 * cos + x1 * sin), -1).to(x.dtype)

    def forward(self, positions, hidden, batch):
        qkv = self.qkv_proj(hidden)                                    # qwen3.py:80
        T = qkv.shape[0]
        q, k, v = qkv.split([self.h * self.hd, self.kv * self.hd, self.kv * self.hd], dim=-1)
        q = self.q_norm(q.reshape(T, self.h, self.hd))                 # :85-87
        k = self.k_norm(k.reshape(T, self.kv, self.hd))
        v = v.reshape(T, self.kv, self.hd)
        q, k = self._rope(positions, q), self._rope(positions, k)      # :88
        S = T // batch
        rep = self.h // self.kv
        qh = q.view(batch, S, self.h, self.hd).transpose(1, 2)
        kh = k.view(batch, S, self.kv, self.hd).transpose(1, 2).repeat_interleave(rep, dim=1)
        vh = v.view(batch, S, self.kv, self.hd).transpose(1, 2).repeat_interleave(rep, dim=1)
        o = F.scaled_dot_product_attention(qh, kh, vh, is_causal=True)  # stands in for :89
        o = o.transpose(1, 2).reshape(T, self.h * self.hd)
        return o


class TinyMLP(nn.Module):
    def __init__(self, cfg, lin, act_cls):
        super().__init__()
        self.gate_up_proj = lin["gate_up"]()
        self.down_proj = lin["down"]()
        self.act_fn = act_cls()


class TinyDecoderLayer(nn.Module):
    def __init__(self, cfg, lin, norm_cls, act_cls):
        super().__init__()
        self.self_attn = TinyAttention(cfg, lin, norm_cls)
        self.mlp = TinyMLP(cfg, lin, act_cls)
        self.input_layernorm = norm_cls(cfg["hidden"], eps=cfg["eps"])
        self.post_attention_layernorm = norm_cls(cfg["hidden"], eps=cfg["eps"])


class TinyModel(nn.Module):
    def __init__(self, cfg, n_layers, vocab, lin, norm_cls, act_cls):
        super().__init__()
        self.embed_tokens = nn.Embedding(vocab, cfg["hidden"])
        self.layers = nn.ModuleList([TinyDecoderLayer(cfg, lin, norm_cls, act_cls) for _ in range(n_layers)])
        self.norm = norm_cls(cfg["hidden"], eps=cfg["eps"])


class TorchRMSNorm(nn.Module):
    """Eager torch form of the reference RMSNorm (layernorm.py:16-40); only used where the model has to
    run without a GPU (calibrating the synthetic corpus on the build container)."""

    def __init__(self, hidden_size, eps=1e-6):
        super().__init__()
        self.eps = eps
        self.weight = nn.Parameter(torch.ones(hidden_size))

    def forward(self, x, residual=None):
        h, res = orc.ref_rmsnorm(x, self.weight.data, self.eps, residual)
        return h if residual is None else (h, res)


class TorchSiluAndMul(nn.Module):
    def forward(self, x):
        return orc.ref_silu_and_mul(x)


class TinyQwen3ForCausalLM(nn.Module):
    """``fused=True`` runs the fused-producer path of the engine: RMSNorm.forward_quant /
    SiluAndMul.forward_quant hand ``QuantizedActivation`` to the quantized TP linears."""

    def __init__(self, n_layers=4, vocab=2048, linear_factory=tp_linears, cfg=CFG_06B, seed=0,
                 device="cuda", ops="repo", lm_head_std=0.12):
        super().__init__()
        if ops == "repo":
            from nanovllm.layers.activation import SiluAndMul
            from nanovllm.layers.layernorm import RMSNorm
            norm_cls, act_cls = RMSNorm, SiluAndMul
        else:
            norm_cls, act_cls = TorchRMSNorm, TorchSiluAndMul
        self.cfg, self.vocab = cfg, vocab
        prev = torch.get_default_dtype()
        torch.set_default_dtype(torch.bfloat16)
        try:
            with torch.device(device):
                self.model = TinyModel(cfg, n_layers, vocab, linear_factory(cfg), norm_cls, act_cls)
                self.lm_head = nn.Linear(cfg["hidden"], vocab, bias=False)
        finally:
            torch.set_default_dtype(prev)
        # identical weights for every variant: one CPU generator, fixed visiting order
        g = torch.Generator().manual_seed(seed)

        def fill(p, std, mean=0.0):
            p.data.copy_((torch.randn(p.shape, generator=g) * std + mean).to(torch.bfloat16))

        fill(self.model.embed_tokens.weight, 1.0)
        for layer in self.model.layers:
            for lin in (layer.self_attn.qkv_proj, layer.self_attn.o_proj, layer.mlp.gate_up_proj,
                        layer.mlp.down_proj):
                fill(lin.weight, 0.03)
            for nrm in (layer.self_attn.q_norm, layer.self_attn.k_norm, layer.input_layernorm,
                        layer.post_attention_layernorm):
                fill(nrm.weight, 0.1, 1.0)
        fill(self.model.norm.weight, 0.1, 1.0)
        fill(self.lm_head.weight, lm_head_std)

    def bf16_weights(self):
        """[(layer, name, weight bf16 CPU)] of the 4 projections, before any quantisation."""
        out = []
        for i, layer in enumerate(self.model.layers):
            for parent, name in (("self_attn", "qkv_proj"), ("self_attn", "o_proj"),
                                 ("mlp", "gate_up_proj"), ("mlp", "down_proj")):
                out.append((i, parent, name, getattr(getattr(layer, parent), name).weight.data.cpu().clone()))
        return out

    def swap_linears(self, make):
        """Replace every projection by ``make(weight_bf16_cpu)`` (used to install OracleLinear)."""
        for i, parent, name, w in self.bf16_weights():
            setattr(getattr(self.model.layers[i], parent), name, make(w))

    @torch.inference_mode()
    def forward(self, input_ids: torch.Tensor, fused: bool = False, qdtype=None) -> torch.Tensor:
        """input_ids [B, S] -> logits fp32 [B*S, vocab].  Token-major 2-D activations [T, hidden]
        through the projections, as the engine feeds them (qwen3.py:147-162, 186-189)."""
        B, S = input_ids.shape
        positions = torch.arange(S, device=input_ids.device).repeat(B)
        h = self.model.embed_tokens(input_ids.reshape(-1))
        residual = None
        for layer in self.model.layers:
            att, mlp = layer.self_attn, layer.mlp
            if fused:
                if residual is None:
                    hq, residual = layer.input_layernorm.forward_quant(h, None, qdtype), h
                else:
                    hq, residual = layer.input_layernorm.forward_quant(h, residual, qdtype)
                h = att.o_proj(att(positions, hq, B))
                hq, residual = layer.post_attention_layernorm.forward_quant(h, residual, qdtype)
                h = mlp.down_proj(mlp.act_fn.forward_quant(mlp.gate_up_proj(hq), qdtype))
            else:
                if residual is None:                                   # qwen3.py:153-156
                    h, residual = layer.input_layernorm(h), h
                else:
                    h, residual = layer.input_layernorm(h, residual)
                h = att.o_proj(att(positions, h, B))                   # :157, :90
                h, residual = layer.post_attention_layernorm(h, residual)   # :160
                h = mlp.down_proj(mlp.act_fn(mlp.gate_up_proj(h)))     # :116-118
        h, _ = self.model.norm(h, residual)                            # :188
        return self.lm_head(h).float()


def synthetic_corpus(teacher_logits: torch.Tensor, seed: int = 1) -> torch.Tensor:
    """Targets drawn from the bf16 teacher's own next-token distribution (seeded): the
    stand-in for WikiText-2 in test_ppl.py:33-68, so that perplexity is a meaningful number (the
    teacher's PPL equals exp(its entropy)) and quantisation noise shows up as a PPL increase."""
    g = torch.Generator().manual_seed(seed)
    p = torch.softmax(teacher_logits.cpu().double(), -1)
    return torch.multinomial(p, 1, generator=g).squeeze(1)


def pseudo_ppl_sampled(logits: torch.Tensor, targets: torch.Tensor) -> float:
    """exp(mean cross-entropy) on ONE sampled corpus, the loop body of test_ppl.py:60-68.  With a
    few hundred tokens its sampling noise (about 1 %) exceeds the 0.5 % bar; reported, not asserted."""
    loss = F.cross_entropy(logits.cpu().double(), targets, reduction="sum")
    return float(torch.exp(loss / targets.numel()))


def pseudo_ppl(logits: torch.Tensor, teacher_logits: torch.Tensor) -> float:
    """Perplexity in expectation over the synthetic corpus: exp of the mean cross-entropy between
    the teacher's next-token distribution and the model's, i.e. the limit of
    ``pseudo_ppl_sampled`` over infinitely many corpora drawn by ``synthetic_corpus``.  Same
    quantity as test_ppl.py's perplexity with the sampling noise of a 256-token corpus removed,
    which is what makes a 0.5 % bar decidable at this size."""
    p = torch.softmax(teacher_logits.cpu().double(), -1)
    ce = -(p * torch.log_softmax(logits.cpu().double(), -1)).sum(-1).mean()
    return float(torch.exp(ce))
