"""Talker LLM (models/qwen3_tts_talker.py:33-136 of the reference): ``model(input_embeds, positions)
-> hidden`` over the decoder stack, ``compute_logits(hidden)`` through ``codec_head`` (last token of
every sequence in prefill), plus the embeddings / ``text_projection`` used for prompt assembly."""
from __future__ import annotations

import torch
from torch import nn

from .. import _native as nat
from ..config import StackConfig
from ..layers.embed_head import ParallelLMHead
from ..layers.linear import ColumnParallelLinear, RowParallelLinear
from .qwen3_tts_share import DecoderStack, load_reference_weights


class Qwen3TTSTalkerResizeMLP(nn.Module):
    """text_hidden -> text_hidden (SiLU) -> hidden, with biases (qwen3_tts_talker.py:14-31)."""

    def __init__(self, input_size: int, intermediate_size: int, output_size: int, bias: bool = True):
        super().__init__()
        self.linear_fc1 = ColumnParallelLinear(input_size, intermediate_size, bias=bias)
        self.linear_fc2 = RowParallelLinear(intermediate_size, output_size, bias=bias)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.linear_fc2(nat.silu_and_mul(self.linear_fc1(x), None))


class Qwen3TTSTalkerModel(DecoderStack):
    def __init__(self, config: StackConfig, rounding: int = nat.ROUND_EAGER):
        super().__init__(config, rounding)
        self.codec_embedding = nn.Embedding(config.vocab_size, config.hidden_size)
        if config.text_vocab_size:
            self.text_embedding = nn.Embedding(config.text_vocab_size, config.text_hidden_size)

    def get_input_embeddings(self):
        return self.codec_embedding

    def get_text_embeddings(self):
        return self.text_embedding

    def forward(self, input_embeds: torch.Tensor, positions: torch.Tensor, tail_weight=None) -> torch.Tensor:
        return self.run_layers(input_embeds, positions, None, tail_weight)


class Qwen3TTSTalkerForCausalLM(nn.Module):
    def __init__(self, config: StackConfig, rounding: int = nat.ROUND_EAGER):
        super().__init__()
        self.config = config
        self.model = Qwen3TTSTalkerModel(config, rounding)
        self.codec_head = ParallelLMHead(config.vocab_size, config.hidden_size)
        if config.text_vocab_size:
            self.text_projection = Qwen3TTSTalkerResizeMLP(config.text_hidden_size, config.text_hidden_size,
                                                           config.hidden_size, bias=True)

    @staticmethod
    def convert_state_dict(state_dict: dict) -> dict:
        """Checkpoint keys -> module keys: drop ``talker.``; skip the predictor and speaker encoder
        (same selection as qwen3_tts_talker.py:81-109)."""
        out = {}
        for key, value in state_dict.items():
            if key.startswith(("talker.code_predictor.", "speaker_encoder.")):
                continue
            out[key[7:] if key.startswith("talker.") else key] = value
        return out

    def load_state_dict(self, state_dict, strict=True):
        return load_reference_weights(self, self.convert_state_dict(state_dict), strict=strict)

    def get_input_embeddings(self):
        return self.model.get_input_embeddings()

    def get_text_embeddings(self):
        return self.model.get_text_embeddings()

    def forward(self, input_embeds: torch.Tensor, positions: torch.Tensor) -> torch.Tensor:
        return self.model(input_embeds, positions, self.codec_head.weight)

    def compute_logits(self, hidden_states: torch.Tensor) -> torch.Tensor:
        return self.codec_head(hidden_states)
