"""Process-global batch-layout context: how the model runner tells the attention layers and the
LM head about the step's layout without threading it through every forward (same names as
utils/context.py:6-27 of the reference).  One addition: ``max_context_len`` (capacity of the block
table in tokens) which bounds the split-KV grid of the decode attention kernel."""
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
    max_context_len: int = 0


_CONTEXT = Context()


def get_context() -> Context:
    return _CONTEXT


def set_context(is_prefill, cu_seqlens_q=None, cu_seqlens_k=None, max_seqlen_q=0, max_seqlen_k=0,
                slot_mapping=None, context_lens=None, block_tables=None, max_context_len=0):
    global _CONTEXT
    _CONTEXT = Context(is_prefill, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, slot_mapping,
                       context_lens, block_tables, max_context_len)


def reset_context():
    global _CONTEXT
    _CONTEXT = Context()
