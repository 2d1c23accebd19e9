"""RoPE modules with the reference's surface (layers/rotary_embedding.py:87-237):
``get_rope(head_size, rotary_dim, max_position, base, rope_scaling)`` -> module with buffer
``inv_freq`` and ``forward(positions, query, key) -> (query, key)``.

cos/sin are recomputed from ``positions`` inside the kernel (fp32 angle -> bf16 cos/sin) and the
rotation keeps the three bf16 roundings of the reference's eager expression.  The talker's 3-D
M-RoPE is a per-column choice of position row (``modality``); with the 1-D positions the runner
feeds (talker_mode_runner.py:134) it collapses to plain RoPE."""
import torch
from torch import nn

from .. import _native as nat


def mrope_dim_modality(head_dim: int, mrope_section, interleaved: bool) -> list[int]:
    """Position row (0..2) read by each of the ``head_dim`` cos/sin columns
    (apply_multimodal_rotary_pos_emb, rotary_embedding.py:52-80)."""
    half = head_dim // 2
    if interleaved:
        m = [0] * half
        n_mod = len(mrope_section)
        for i, n in enumerate(mrope_section[1:], 1):
            for f in range(i, min(n * n_mod, half), n_mod):
                m[f] = i
        return m + m
    out = []
    for i, n in enumerate(mrope_section):          # sections doubled in WIDTH, rotary_embedding.py:74
        out += [i % 3] * (2 * n)
    assert len(out) == head_dim, (len(out), head_dim)
    return out


class RotaryEmbedding(nn.Module):
    def __init__(self, head_size: int, rotary_dim: int, max_position_embeddings: int, base: float) -> None:
        super().__init__()
        assert rotary_dim == head_size
        self.head_size = head_size
        inv_freq = 1.0 / (base ** (torch.arange(0, rotary_dim, 2, dtype=torch.float) / rotary_dim))
        self.register_buffer("inv_freq", inv_freq, persistent=False)
        self.attention_scaling = 1.0
        self.register_buffer("modality", None, persistent=False)

    def forward(self, positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor):
        """query [N, heads*hd] or [N, heads, hd]; key likewise.  Returns tensors of the same shape."""
        hd = self.head_size
        q3, k3 = query.view(query.shape[0], -1, hd), key.view(key.shape[0], -1, hd)
        pos = positions
        if pos.dim() == 3:                      # [3, 1, N] as TalkerRotaryEmbedding accepts
            pos = pos[:, 0]
        q, k = nat.qknorm_rope_kvwrite(q3, k3, None, None, None, 0.0, pos, self.inv_freq, modality=self.modality,
                                       attention_scaling=self.attention_scaling)
        return q.view(query.shape), k.view(key.shape)


class TalkerRotaryEmbedding(RotaryEmbedding):
    def __init__(self, head_size, rotary_dim, max_position_embeddings, base, mrope_section, mrope_interleaved=False):
        super().__init__(head_size, rotary_dim, max_position_embeddings, base)
        self.mrope_section = list(mrope_section)
        self.mrope_interleaved = mrope_interleaved
        mod = torch.tensor(mrope_dim_modality(head_size, self.mrope_section, mrope_interleaved), dtype=torch.int32)
        self.modality = mod


def get_rope(head_size: int, rotary_dim: int, max_position: int, base: float, rope_scaling: dict | None = None):
    if rope_scaling and "mrope_section" in rope_scaling:
        return TalkerRotaryEmbedding(head_size, rotary_dim, max_position, base, rope_scaling["mrope_section"],
                                     rope_scaling.get("interleaved", rope_scaling.get("mrope_interleaved", False)))
    return RotaryEmbedding(head_size, rotary_dim, max_position, base)
