"""Prefill-first continuous batching (engine/scheduler.py:8-79, talker_llm_engine.py:11-83,
predictor_llm_engine.py:15-30 of the reference): waiting sequences are prefilled as one packed
batch, otherwise every running sequence decodes one token.  The talker variant only schedules
sequences whose next input embedding has been fed."""
from __future__ import annotations

from collections import deque

from ..config import Config
from .block_manager import BlockManager
from .sequence import Sequence, SequenceStatus


class Scheduler:
    def __init__(self, config: Config):
        self.max_num_seqs = config.max_num_seqs
        self.max_num_batched_tokens = config.max_num_batched_tokens
        self.eos = config.eos
        self.block_manager = BlockManager(config.num_kvcache_blocks, config.kvcache_block_size)
        self.waiting: deque[Sequence] = deque()
        self.running: deque[Sequence] = deque()
        self.request_id_to_seq: dict[str, Sequence] = {}

    def is_finished(self):
        return not self.waiting and not self.running

    def add(self, seq: Sequence):
        self.waiting.append(seq)

    def _schedule_prefill(self) -> list[Sequence]:
        picked, tokens = [], 0
        while self.waiting and len(picked) < self.max_num_seqs:
            seq = self.waiting[0]
            if tokens + len(seq) > self.max_num_batched_tokens or not self.block_manager.can_allocate(seq):
                break
            self.block_manager.allocate(seq)
            tokens += len(seq) - seq.num_cached_tokens
            seq.status = SequenceStatus.RUNNING
            self.waiting.popleft()
            self.running.append(seq)
            picked.append(seq)
        return picked

    def _ready(self, seq: Sequence) -> bool:
        return True

    def _schedule_decode(self) -> list[Sequence]:
        # fast path (every steady-state decode step): all running sequences fit, all are ready and none opens a new
        # KV page this step, so the batch is the running queue in order and the queue is left as it is - exactly what
        # the general loop below produces, without its per-sequence pops and allocator calls
        n = len(self.running)
        if 0 < n <= self.max_num_seqs:
            bs, ready = self.block_manager.block_size, self._ready
            if all(s.num_tokens % bs != 1 and ready(s) for s in self.running):
                return list(self.running)
        picked = []
        for _ in range(len(self.running)):
            if not self.running or len(picked) >= self.max_num_seqs:
                break
            seq = self.running.popleft()
            if not self._ready(seq):
                self.running.append(seq)
                continue
            while not self.block_manager.can_append(seq):
                if self.running:
                    self.preempt(self.running.pop())
                else:
                    self.preempt(seq)
                    break
            else:
                self.block_manager.may_append(seq)
                picked.append(seq)
        self.running.extendleft(reversed(picked))
        return picked

    def schedule(self) -> tuple[list[Sequence], bool]:
        picked = self._schedule_prefill()
        if picked:
            return picked, True
        return self._schedule_decode(), False

    def preempt(self, seq: Sequence):
        """Drop the KV pages and queue the sequence for a full re-prefill (scheduler.py:68-71)."""
        seq.status = SequenceStatus.WAITING
        self.block_manager.deallocate(seq)
        self.waiting.appendleft(seq)

    def _finish(self, seq: Sequence):
        seq.status = SequenceStatus.FINISHED
        if seq.request_id is not None:
            self.request_id_to_seq.pop(seq.request_id, None)
        self.block_manager.deallocate(seq)
        self.running.remove(seq)

    def clear_request(self, request_id: str):
        seq = self.request_id_to_seq.pop(request_id, None)
        if seq is None:
            return
        if seq.block_table:
            self.block_manager.deallocate(seq)
        if seq in self.running:
            self.running.remove(seq)
        if seq in self.waiting:
            self.waiting.remove(seq)

    def postprocess(self, seqs: list[Sequence], token_ids: list[int]):
        for seq, tok in zip(seqs, token_ids):
            seq.append_token(tok, None)
            if (not seq.ignore_eos and tok == self.eos) or seq.num_completion_tokens >= seq.max_tokens:
                self._finish(seq)


class TalkerScheduler(Scheduler):
    def _ready(self, seq: Sequence) -> bool:
        return not (len(seq) > 0 and seq.decode_input_embeds is None)

    def postprocess(self, seqs, token_ids, hidden_states):
        rows = hidden_states.unbind(0)             # the row views in one call (indexing per sequence costs ~1.5 us each)
        for i, (seq, tok) in enumerate(zip(seqs, token_ids)):
            seq.append_token(tok, rows[i])
            seq.decode_input_embeds = None
            seq.batch_src = None
            hit_eos = not seq.ignore_eos and tok == self.eos
            # requests driven by the interface stop on the codec EOS there, not here
            # (talker_llm_engine.py:74-77)
            if hit_eos or (seq.request_id is None and seq.num_completion_tokens >= seq.max_tokens):
                self._finish(seq)


class PredictorScheduler(Scheduler):
    pass
