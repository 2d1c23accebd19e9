"""Per-request state read by the scheduler and the model runner; same field names as the
reference's ``Sequence`` (engine/sequence.py:15-90) because the runner contract
is expressed in them.  Sequences are embed-only: the prompt arrives as ``input_embeds [1, T, H]``
and ``token_ids`` holds generated codec ids only."""
from __future__ import annotations

from enum import Enum, auto
from itertools import count

import torch

from ..sampling_params import SamplingParams


class SequenceStatus(Enum):
    WAITING = auto()
    RUNNING = auto()
    FINISHED = auto()


class Sequence:
    block_size = 256
    counter = count()

    def __init__(self, token_ids: list[int] | None, sampling_params: SamplingParams | None = None,
                 input_embeds: torch.Tensor | None = None, request_id: str | None = None):
        sp = sampling_params or SamplingParams()
        self.seq_id = next(Sequence.counter)
        self.status = SequenceStatus.WAITING
        self.input_embeds = input_embeds
        self.request_id = request_id
        self.decode_input_embeds: torch.Tensor | None = None
        self.token_ids = list(token_ids or [])
        self.last_token = self.token_ids[-1] if self.token_ids else None
        self.num_prompt_tokens = len(self.token_ids)
        self.num_tokens = len(self.token_ids) if self.token_ids else int(input_embeds.shape[1])
        self.num_cached_tokens = 0
        self.block_table: list[int] = []
        self.temperature = sp.temperature
        self.max_tokens = sp.max_tokens
        self.ignore_eos = sp.ignore_eos
        self.top_k = sp.top_k
        self.generation_steps = 0
        self.last_hidden_state = None

    def __len__(self):
        return self.num_tokens

    @property
    def is_finished(self):
        return self.status == SequenceStatus.FINISHED

    @property
    def num_completion_tokens(self):
        return self.num_tokens - self.num_prompt_tokens

    @property
    def completion_token_ids(self):
        return self.token_ids[self.num_prompt_tokens:]

    @property
    def num_cached_blocks(self):
        return self.num_cached_tokens // self.block_size

    @property
    def num_blocks(self):
        return (self.num_tokens + self.block_size - 1) // self.block_size

    @property
    def last_block_num_tokens(self):
        return self.num_tokens - (self.num_blocks - 1) * self.block_size

    def append_token(self, token_id: int, last_hidden_state=None):
        self.token_ids.append(token_id)
        self.last_token = token_id
        self.num_tokens += 1
        self.last_hidden_state = last_hidden_state


class PredictorSequence(Sequence):
    """predictor_model_runner.py:25-28: adds the codebook index being generated."""

    def __init__(self, token_ids, sampling_params=None, input_embeds=None, generation_steps: int = 0, request_id=None):
        super().__init__(token_ids, sampling_params, input_embeds, request_id)
        self.generation_steps = generation_steps
