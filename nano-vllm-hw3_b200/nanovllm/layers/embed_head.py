"""Quantized LM head.

Mirrors ``ParallelLMHead`` of the reference's nanovllm/layers/embed_head.py:44-66 (constructor,
vocab-parallel ``weight`` shard ``[vocab/tp, hidden]``, ``weight_loader``, rank-0 gather of the
logits) and puts the W8A8 path inside it: ``quantize_()`` turns the shard into (``qw``, ``sw``) and
``forward`` runs ``ql_linear``.  vocab x hidden is the largest matrix of a decode step (151 936 x
1024 = 311 MB in bf16 on Qwen3-0.6B, 156 MB in int8 -- versus 440 MB for all 112 layer matrices),
so it is the single largest HBM stream the path can halve.

The embedding lookup itself (``VocabParallelEmbedding.forward``, :35-42) is out of scope; with tied
embeddings (Qwen3-0.6B) pass ``keep_weight=True`` so the bf16 table stays available to it.
"""
from __future__ import annotations

import torch
import torch.distributed as dist
import torch.nn.functional as F
from torch import nn

from nanovllm import _ext
from nanovllm.utils.context import get_context


def _tp_rank_size():
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


class ParallelLMHead(nn.Module):

    def __init__(self, num_embeddings: int, embedding_dim: int, bias: bool = False):
        assert not bias                                           # reference :52
        super().__init__()
        self.tp_rank, self.tp_size = _tp_rank_size()
        assert num_embeddings % self.tp_size == 0                 # reference :19
        self.num_embeddings = num_embeddings
        self.num_embeddings_per_partition = num_embeddings // self.tp_size
        self.vocab_start_idx = self.num_embeddings_per_partition * self.tp_rank
        self.vocab_end_idx = self.vocab_start_idx + self.num_embeddings_per_partition
        self.weight = nn.Parameter(torch.empty(self.num_embeddings_per_partition, embedding_dim))
        self.weight.weight_loader = self.weight_loader
        self.qdtype = None
        self.qw = self.sw = None

    def weight_loader(self, param: nn.Parameter, loaded_weight: torch.Tensor):
        shard = param.data.size(0)                                # reference :27-32
        param.data.copy_(loaded_weight.narrow(0, self.tp_rank * shard, shard))

    @torch.no_grad()
    def quantize_(self, dtype: torch.dtype, keep_weight: bool = False) -> "ParallelLMHead":
        """Per-row (per vocabulary entry) quantisation of this rank's shard."""
        from nanovllm.utils import quantization as Q
        assert dtype in [torch.float8_e4m3fn, torch.int8]
        self.qw, self.sw = Q.per_row_quant(self.weight.data, dtype)
        self.qdtype = dtype
        if not keep_weight:
            self.weight = None
        return self

    def forward(self, x: torch.Tensor):
        """``x`` bf16 [T, hidden] -> logits (rank 0; None elsewhere under TP) -- the reference's
        signature and behaviour, embed_head.py:56-66: during prefill (``get_context().is_prefill``)
        only the last token of every sequence (``cu_seqlens_q[1:] - 1``) is projected."""
        context = get_context()
        if context.is_prefill:
            last_indices = context.cu_seqlens_q[1:] - 1
            x = x[last_indices].contiguous()
        if self.qw is None:
            logits = F.linear(x, self.weight)
        else:
            logits = _ext.linear(x, self.qw, self.sw, None, self.qdtype, 0)
        if self.tp_size > 1:
            all_logits = [torch.empty_like(logits) for _ in range(self.tp_size)] if self.tp_rank == 0 else None
            dist.gather(logits, all_logits, 0)
            logits = torch.cat(all_logits, -1) if self.tp_rank == 0 else None
        return logits
