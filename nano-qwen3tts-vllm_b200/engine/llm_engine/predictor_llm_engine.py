"""``PredictorLLMEngine`` step API (reference engine/llm_engine/predictor_llm_engine.py:33-66)."""
from __future__ import annotations

import torch

from ...sampling_params import SamplingParams
from ..model_runner.predictor_model_runner import PredictorModelRunner
from ..scheduler import PredictorScheduler
from ..sequence import PredictorSequence
from .base import LLMEngine


class PredictorLLMEngine(LLMEngine):
    runner_cls = PredictorModelRunner
    scheduler_cls = PredictorScheduler

    def add_request(self, inputs_embeds: list[torch.Tensor], sampling_params: SamplingParams | list[SamplingParams],
                    request_id: str | None = None):
        if not isinstance(sampling_params, list):
            sampling_params = [sampling_params] * len(inputs_embeds)
        for emb, sp in zip(inputs_embeds, sampling_params):
            seq = PredictorSequence([], input_embeds=self._to_device(emb), sampling_params=sp, generation_steps=0,
                                    request_id=request_id)
            if request_id is not None:
                self.scheduler.request_id_to_seq[request_id] = seq
            self.scheduler.add(seq)

    def step(self):
        seqs, is_prefill = self.scheduler.schedule()
        if not seqs:
            return [], 0
        token_ids = self.model_runner.call("run", seqs, is_prefill)
        self.scheduler.postprocess(seqs, token_ids)
        for s in seqs:
            s.generation_steps += 1
        outputs = [(s.request_id, s.seq_id, s.completion_token_ids) for s in seqs if s.is_finished]
        return outputs, (sum(len(s) for s in seqs) if is_prefill else -len(seqs))
