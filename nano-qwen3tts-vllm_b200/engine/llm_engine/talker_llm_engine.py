"""``TalkerLLMEngine`` step API (reference engine/llm_engine/talker_llm_engine.py:86-134)."""
from __future__ import annotations

import torch

from ...sampling_params import SamplingParams
from ..model_runner.talker_mode_runner import TalkerModeModelRunner
from ..scheduler import TalkerScheduler
from ..sequence import Sequence
from .base import LLMEngine


class TalkerLLMEngine(LLMEngine):
    runner_cls = TalkerModeModelRunner
    scheduler_cls = TalkerScheduler

    def add_request(self, inputs_embeds: list[torch.Tensor], sampling_params: SamplingParams | list[SamplingParams],
                    request_id: str | None = None):
        """New ``request_id`` -> a prefill of ``inputs_embeds[i] [1, T, H]``; known ``request_id`` ->
        the tensor is the next decode input of that stream."""
        if not isinstance(sampling_params, list):
            sampling_params = [sampling_params] * len(inputs_embeds)
        for emb, sp in zip(inputs_embeds, sampling_params):
            emb = self._to_device(emb)
            if request_id is not None and request_id in self.scheduler.request_id_to_seq:
                seq = self.scheduler.request_id_to_seq[request_id]
                seq.decode_input_embeds = emb
                seq.batch_src = None
                return
            seq = Sequence([], input_embeds=emb, sampling_params=sp, request_id=request_id)
            if request_id is not None:
                self.scheduler.request_id_to_seq[request_id] = seq
            self.scheduler.add(seq)

    def add_requests(self, inputs_embeds: torch.Tensor, sampling_params: SamplingParams, request_ids: list[str | None]):
        """Batched ``add_request`` for decode inputs (and equal-length prompts): ``inputs_embeds`` [n, T, H] (host,
        ideally pinned) crosses PCIe as ONE copy; row i belongs to ``request_ids[i]``."""
        dev = self._to_device(inputs_embeds)
        known = self.scheduler.request_id_to_seq
        rows = dev.unbind(0)                               # n row views from one call
        for i, rid in enumerate(request_ids):
            seq = known.get(rid) if rid is not None else None
            if seq is not None:
                seq.decode_input_embeds = rows[i]
                seq.batch_src = (dev, i)
                continue
            seq = Sequence([], input_embeds=dev[i:i + 1], sampling_params=sampling_params, request_id=rid)
            if rid is not None:
                known[rid] = seq
            self.scheduler.add(seq)

    def step_launch(self):
        """Schedule and QUEUE one step; returns a handle for ``step_collect`` (None: nothing to run).  Between the two
        calls the GPU is busy and the host is free - e.g. to queue another engine's work that only needs this step's
        device results (``model_runner.post_sample_hook``)."""
        seqs, is_prefill = self.scheduler.schedule()
        if not seqs:
            return None
        return seqs, is_prefill, self.model_runner.call("run_launch", seqs, is_prefill)

    def step_collect(self, launched):
        """Wait for a launched step; returns what ``step_with_outputs`` returns."""
        if launched is None:
            return [], 0, []
        seqs, is_prefill, pending = launched
        token_ids, hidden = self.model_runner.call("run_collect", pending)
        self.scheduler.postprocess(seqs, token_ids, hidden)
        self.last_step = (seqs, token_ids, hidden)      # the step's batch as a whole: hidden [n, H] on the device
        num_tokens = sum(len(s) for s in seqs) if is_prefill else -len(seqs)      # counted after the append, as the reference
        outputs = [(s.request_id, s.seq_id, s.completion_token_ids, s.last_hidden_state) for s in seqs if s.is_finished]
        outputs_all = [(s.request_id, s.seq_id, s.completion_token_ids, s.last_hidden_state, s.is_finished)
                       for s in seqs]
        return outputs, num_tokens, outputs_all

    def _step(self):
        seqs, is_prefill = self.scheduler.schedule()
        if not seqs:
            return None, 0, []
        token_ids, hidden = self.model_runner.call("run", seqs, is_prefill)
        self.scheduler.postprocess(seqs, token_ids, hidden)
        self.last_step = (seqs, token_ids, hidden)      # the step's batch as a whole: hidden [n, H] on the device
        return seqs, (sum(len(s) for s in seqs) if is_prefill else -len(seqs)), seqs

    def step(self):
        seqs, num_tokens, _ = self._step()
        if seqs is None:
            return [], 0
        return [(s.request_id, s.seq_id, s.completion_token_ids, s.last_hidden_state) for s in seqs if s.is_finished], \
            num_tokens

    def step_with_outputs(self):
        seqs, num_tokens, _ = self._step()
        if seqs is None:
            return [], 0, []
        outputs = [(s.request_id, s.seq_id, s.completion_token_ids, s.last_hidden_state) for s in seqs if s.is_finished]
        outputs_all = [(s.request_id, s.seq_id, s.completion_token_ids, s.last_hidden_state, s.is_finished)
                       for s in seqs]
        return outputs, num_tokens, outputs_all
