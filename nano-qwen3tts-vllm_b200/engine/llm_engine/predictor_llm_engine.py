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

    def add_requests(self, inputs_embeds: torch.Tensor, sampling_params: SamplingParams, request_ids: list[str | None]):
        """Batched ``add_request``: ``inputs_embeds`` [n, 2, H] (host, ideally pinned) crosses PCIe as ONE copy; request
        i gets row i.  Same queueing as n single calls."""
        dev = self._to_device(inputs_embeds)
        rows = dev.unsqueeze(1).unbind(0)                  # n views [1, 2, H] from one call (a slice per request costs ~1.5 us each)
        known, waiting = self.scheduler.request_id_to_seq, self.scheduler.waiting
        for i, rid in enumerate(request_ids):
            seq = PredictorSequence([], input_embeds=rows[i], sampling_params=sampling_params, generation_steps=0,
                                    request_id=rid)
            seq.batch_src = (dev, i)
            if rid is not None:
                known[rid] = seq
            waiting.append(seq)

    burst = True          # whole-request graphs (PredictorModelRunner.run_burst); False = step by step as the reference

    def _burst_launch(self, seqs):
        """Every scheduled sequence is a fresh 2-row request: one graph replay yields all of its codes.  Queues the
        replay, the caller's hook and the copy of the codes to a pinned buffer; ``_burst_collect`` waits."""
        n_codes = self.model_runner.burst_codes()
        codes = self.model_runner.call("run_burst", seqs)                    # replay queued; the GPU is busy from here
        hook = getattr(self, "burst_hook", None)
        if hook is not None:
            hook(seqs, codes)                    # caller work on the DEVICE codes, queued before the burst's one synchronisation
        n = len(seqs)
        pins = self.__dict__.setdefault("_code_pins", [None, None, 0])
        slot = pins[2] = pins[2] ^ 1
        if pins[slot] is None or pins[slot].shape[0] < n:
            pins[slot] = torch.empty(max(n, 64), n_codes, dtype=torch.int64).pin_memory()
        host = pins[slot][:n]
        host.copy_(codes, non_blocking=True)
        done = torch.cuda.Event()
        done.record()
        for s in seqs:                                                       # bookkeeping that needs no code values
            s.num_tokens += n_codes
            s.generation_steps += n_codes
            self.scheduler._finish(s)
        return seqs, host, done

    def _burst_collect(self, launched):
        seqs, host, done = launched
        done.synchronize()                                                   # the one synchronisation of the burst
        toks = host.tolist()
        for s, row in zip(seqs, toks):
            s.token_ids.extend(row)
            s.last_token = row[-1]
        return [(s.request_id, s.seq_id, s.completion_token_ids) for s in seqs], 2 * len(seqs)

    def _step_burst(self, seqs):
        return self._burst_collect(self._burst_launch(seqs))

    def step_launch(self):
        """Schedule and QUEUE one step without waiting for the GPU when the batch is a whole-request burst (otherwise the
        step runs to completion here); ``step_collect`` returns what ``step`` returns."""
        seqs, is_prefill = self.scheduler.schedule()
        if not seqs:
            return None
        if is_prefill and self.burst and self.scheduler.eos < 0 and self.model_runner.burst_ok(seqs):
            return "burst", self._burst_launch(seqs)
        return "done", self._step_scheduled(seqs, is_prefill)

    def step_collect(self, launched):
        if launched is None:
            return [], 0
        kind, payload = launched
        return self._burst_collect(payload) if kind == "burst" else payload

    # The predictor never stops on an EOS id (config.eos stays -1: no tokenizer, predictor_llm_engine.py:33-66 of the
    # reference finishes on max_tokens), so the sampled codes are only needed on the host when a sequence finishes.
    # With ``defer_tokens`` the per-step ``.tolist()`` of the reference (one PCIe round trip and one stream
    # synchronisation per codebook) becomes one read per burst; results are identical.
    defer_tokens = True
    PENDING = -2                                  # placeholder id in Sequence.token_ids until the codes are read back

    def _resolve(self, seqs):
        """Read every deferred step back in one transfer and patch the placeholders of ``seqs``."""
        steps = getattr(self, "_steps", None)
        if not steps:
            return
        sizes = [int(t.numel()) for t in steps]
        flat = torch.cat(steps).tolist()                                     # the one synchronisation of the burst
        offs = [0]
        for n in sizes:
            offs.append(offs[-1] + n)
        for s in seqs:
            refs = getattr(s, "_deferred", None)
            if not refs:
                continue
            base = len(s.token_ids) - len(refs)
            for k, (step, row) in enumerate(refs):
                s.token_ids[base + k] = flat[offs[step] + row]
            s.last_token = s.token_ids[-1]
            s._deferred = []

    def _step_deferred(self, seqs, is_prefill):
        steps = self.__dict__.setdefault("_steps", [])
        prev = None
        if not is_prefill:
            last = len(steps) - 1
            rows = [s._deferred[-1] if getattr(s, "_deferred", None) else None for s in seqs]
            if all(r is not None and r[0] == last for r in rows):
                idx = [r[1] for r in rows]
                prev = steps[last] if idx == list(range(steps[last].numel())) else \
                    steps[last][torch.tensor(idx, device=steps[last].device)]
            elif any(r is not None for r in rows):
                self._resolve(seqs)                                          # mixed batch: fall back to host tokens
        tokens = self.model_runner.call("run_deferred", seqs, is_prefill, prev)
        steps.append(tokens)
        for row, s in enumerate(seqs):
            if not hasattr(s, "_deferred"):
                s._deferred = []
            s._deferred.append((len(steps) - 1, row))
        self.scheduler.postprocess(seqs, [self.PENDING] * len(seqs))
        finished = [s for s in seqs if s.is_finished]
        if finished:
            self._resolve(finished)
        if not self.scheduler.running and not self.scheduler.waiting:
            steps.clear()                                                    # nothing refers to the old steps any more
        elif len(steps) > 64:                                                # continuous traffic: bound the backlog
            self._resolve(list(self.scheduler.running))
            steps.clear()

    def step(self):
        seqs, is_prefill = self.scheduler.schedule()
        if not seqs:
            return [], 0
        return self._step_scheduled(seqs, is_prefill)

    def _step_scheduled(self, seqs, is_prefill):
        if is_prefill and self.burst and self.scheduler.eos < 0 and self.model_runner.burst_ok(seqs):
            return self._step_burst(seqs)
        if self.defer_tokens and self.scheduler.eos < 0 and all(not s.token_ids or hasattr(s, "_deferred") for s in seqs):
            self._step_deferred(seqs, is_prefill)
            for s in seqs:
                s.generation_steps += 1
            outputs = [(s.request_id, s.seq_id, s.completion_token_ids) for s in seqs if s.is_finished]
            return outputs, (sum(len(s) for s in seqs) if is_prefill else -len(seqs))
        token_ids = self.model_runner.call("run", seqs, is_prefill)
        self.scheduler.postprocess(seqs, token_ids)
        for s in seqs:
            s.generation_steps += 1
        outputs = [(s.request_id, s.seq_id, s.completion_token_ids) for s in seqs if s.is_finished]
        return outputs, (sum(len(s) for s in seqs) if is_prefill else -len(seqs))
