"""``Sampler`` (layers/sampler.py:15-24): temperature -> top-k (ties kept) -> softmax -> one
multinomial draw, as ONE kernel.  The Exp(1) noise torch.multinomial would draw internally
(``empty_like(probs).exponential_(1)``) is drawn here with the same call on the same generator, so
for an identical generator state the sampled indices are bit-identical to the reference's."""
import torch
from torch import nn

from .. import _native as nat


class Sampler(nn.Module):
    def forward(self, logits: torch.Tensor, temperatures: torch.Tensor, top_k: int = 50, top_p: float = 1.0,
                noise: torch.Tensor | None = None, out: torch.Tensor | None = None) -> torch.Tensor:
        # top_p is accepted and ignored, exactly like the reference (sampler.py:15-22)
        if noise is None:
            noise = torch.empty(logits.shape, dtype=torch.float32, device=logits.device).exponential_(1)
        return nat.topk_sample(logits, temperatures, noise, top_k=top_k, out=out)
