"""``Silu`` (layers/activation.py:7-13) and the gated product of the MLP
(models/qwen3_tts_share.py:131-135) as one kernel over a fused gate_up tensor."""
import torch
from torch import nn

from .. import _native as nat


class Silu(nn.Module):
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return nat.silu_and_mul(x, None)


class SiluAndMul(nn.Module):
    """x = [gate | up] along the last dim (the output of a merged gate_up projection)."""

    def __init__(self, rounding: int = nat.ROUND_EAGER):
        super().__init__()
        self.rounding = rounding

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        inter = x.shape[-1] // 2
        x2 = x.view(-1, x.shape[-1])
        out = nat.silu_and_mul(x2[:, :inter], x2[:, inter:], self.rounding)
        return out.view(*x.shape[:-1], inter)
