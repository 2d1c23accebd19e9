"""``Qwen3TTSRMSNorm`` (layers/layernorm.py:7-53 of the reference) on libb200tts.

``forward(x)`` -> y;  ``forward(x, residual)`` -> (y, residual + x) with the reference's
add_rms_forward semantics (fp32 sum feeds the norm, layernorm.py:35-38)."""
import torch
from torch import nn

from .. import _native as nat


class Qwen3TTSRMSNorm(nn.Module):
    def __init__(self, hidden_size: int, eps: float = 1e-6, rounding: int = nat.ROUND_EAGER) -> None:
        super().__init__()
        self.eps = eps
        self.rounding = rounding
        self.weight = nn.Parameter(torch.ones(hidden_size), requires_grad=False)

    def rms_forward(self, x: torch.Tensor) -> torch.Tensor:
        return nat.rmsnorm(x, self.weight, self.eps, self.rounding)

    def add_rms_forward(self, x: torch.Tensor, residual: torch.Tensor):
        return nat.add_rmsnorm(x, residual, self.weight, self.eps, self.rounding, round_sum_first=False)

    def forward(self, x: torch.Tensor, residual: torch.Tensor | None = None):
        if residual is None:
            return self.rms_forward(x)
        return self.add_rms_forward(x, residual)
