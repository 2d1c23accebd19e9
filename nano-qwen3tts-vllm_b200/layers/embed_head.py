"""``ParallelLMHead`` (layers/embed_head.py:45-66): in prefill only the last token of every
sequence (``cu_seqlens_q[1:] - 1``) reaches the head GEMM."""
import torch
from torch import nn

from .. import _native as nat
from ..utils.context import get_context
from .linear import linear


class ParallelLMHead(nn.Module):
    def __init__(self, num_embeddings: int, embedding_dim: int, bias: bool = False):
        super().__init__()
        assert not bias
        self.num_embeddings = num_embeddings
        self.embedding_dim = embedding_dim
        self.weight = nn.Parameter(torch.empty(num_embeddings, embedding_dim), requires_grad=False)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        ctx = get_context()
        if ctx.is_prefill:
            last = (ctx.cu_seqlens_q[1:] - 1).to(torch.int64)
            x = nat.gather_rows(x, last)
        return linear(x, self.weight)
