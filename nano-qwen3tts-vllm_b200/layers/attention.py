"""``store_kvcache`` and ``Attention`` with the reference's surface (layers/attention.py:33-96):
``Attention(num_heads, head_dim, scale, num_kv_heads)`` owns ``k_cache`` / ``v_cache`` (empty until
the model runner binds views of the paged pool, base.py:160-164) and ``forward(q, k, v)`` reads the
batch layout from the global Context."""
import torch
from torch import nn

from .. import _native as nat
from ..utils.context import get_context


def store_kvcache(key: torch.Tensor, value: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                  slot_mapping: torch.Tensor) -> None:
    n, num_heads, head_dim = key.shape
    d = num_heads * head_dim
    assert key.stride(-1) == 1 and value.stride(-1) == 1
    assert key.stride(1) == head_dim and value.stride(1) == head_dim
    assert k_cache.stride(1) == d and v_cache.stride(1) == d
    assert slot_mapping.numel() == n
    nat.store_kvcache(key, value, k_cache, v_cache, slot_mapping)


class Attention(nn.Module):
    def __init__(self, num_heads, head_dim, scale, num_kv_heads):
        super().__init__()
        self.num_heads, self.head_dim, self.scale, self.num_kv_heads = num_heads, head_dim, scale, num_kv_heads
        self.k_cache = self.v_cache = torch.tensor([])
        self.workspace = None                     # split-KV scratch, bound by the runner
        self.fused_workspace = None               # zero-filled scratch of the fused decode kernel, bound by the runner

    def attend(self, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, out: torch.Tensor | None = None):
        """Attention only (the KV rows of this step are already in the cache)."""
        ctx = get_context()
        if ctx.is_prefill:
            if ctx.block_tables is not None:      # prefix-cache hit: read K/V through the block table
                return nat.varlen_attn_prefill(q, self.k_cache, self.v_cache, ctx.cu_seqlens_q, ctx.cu_seqlens_k,
                                               ctx.max_seqlen_q, ctx.max_seqlen_k, self.scale,
                                               block_table=ctx.block_tables, out=out)
            return nat.varlen_attn_prefill(q, k, v, ctx.cu_seqlens_q, ctx.cu_seqlens_k, ctx.max_seqlen_q,
                                           ctx.max_seqlen_k, self.scale, out=out)
        max_ctx = ctx.max_context_len or ctx.block_tables.shape[1] * self.k_cache.shape[1]
        return nat.paged_attn_decode(q, self.k_cache, self.v_cache, ctx.block_tables, ctx.context_lens, self.scale,
                                     max_ctx, out=out, workspace=self.workspace)

    def forward(self, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor):
        ctx = get_context()
        if self.k_cache.numel() and self.v_cache.numel():
            store_kvcache(k, v, self.k_cache, self.v_cache, ctx.slot_mapping)
        return self.attend(q, k, v)
