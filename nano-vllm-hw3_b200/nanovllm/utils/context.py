"""Forward context read by ``ParallelLMHead.forward`` -- the interface of the reference's
nanovllm/utils/context.py:5-27 (``Context`` fields, ``get_context`` / ``set_context`` /
``reset_context``), kept so that the quantized LM head is called exactly like the reference's
(``lm_head(hidden_states)``, models/qwen3.py:212-216): during prefill the engine publishes
``cu_seqlens_q`` here and the head keeps the last token of every sequence (embed_head.py:56-59).
Only the fields the hot path's caller contract needs are interpreted by this package.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch


@dataclass
class Context:
    is_prefill: bool = False
    cu_seqlens_q: torch.Tensor | None = None
    cu_seqlens_k: torch.Tensor | None = None
    max_seqlen_q: int = 0
    max_seqlen_k: int = 0
    slot_mapping: torch.Tensor | None = None
    context_lens: torch.Tensor | None = None
    block_tables: torch.Tensor | None = None


_CONTEXT = Context()


def get_context() -> Context:
    return _CONTEXT


def set_context(is_prefill, cu_seqlens_q=None, cu_seqlens_k=None, max_seqlen_q=0, max_seqlen_k=0,
                slot_mapping=None, context_lens=None, block_tables=None) -> None:
    global _CONTEXT
    _CONTEXT = Context(is_prefill, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k,
                       slot_mapping, context_lens, block_tables)


def reset_context() -> None:
    global _CONTEXT
    _CONTEXT = Context()
