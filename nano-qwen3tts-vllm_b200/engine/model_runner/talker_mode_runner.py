"""Talker runner (reference engine/model_runner/talker_mode_runner.py:23-234): decode inputs are
embeddings fed by the frame loop, ``run`` returns ``(token_ids, hidden_states)``.

Quirk kept on purpose: the decode RoPE position is ``len(seq)`` while the KV slot
index is ``len(seq) - 1`` (talker_mode_runner.py:126-128)."""
from __future__ import annotations

import torch

from ...models.qwen3_tts_talker import Qwen3TTSTalkerForCausalLM
from ...utils.context import reset_context
from ..sequence import Sequence
from .base import BF16, ModelRunner


class TalkerModeModelRunner(ModelRunner):
    is_talker = True

    def pick_stack_config(self, model_config):
        return model_config.talker

    def build_model(self):
        return Qwen3TTSTalkerForCausalLM(self.stack_config, self.rounding)

    def graph_batch_sizes(self):
        if self.config.graph_batch_sizes:
            return sorted(self.config.graph_batch_sizes)
        return [1, 2, 4, 8] + list(range(16, self.max_bs + 1, 16))          # talker_mode_runner.py:169

    def post_init(self):
        self.static_embeds = torch.zeros(min(self.config.max_num_seqs, 512), self.stack_config.hidden_size,
                                         dtype=BF16, device=torch.device(self.config.device))
        super().post_init()

    def decode_position(self, seq: Sequence) -> int:
        return len(seq)                                                      # talker_mode_runner.py:126

    def decode_body(self, rows: int) -> torch.Tensor:
        return self.model(self.static_embeds[:rows], self.layout.d["positions"][:rows])

    @torch.inference_mode()
    def run_model(self, seqs: list[Sequence], is_prefill: bool):
        n = len(seqs)
        if is_prefill:
            embeds, positions = self.prepare_prefill(seqs)
            hidden = self.model(embeds, positions)
            logits = self.model.compute_logits(hidden)                       # last token per sequence
            from ...utils.context import get_context
            last = (get_context().cu_seqlens_q[1:] - 1).to(torch.int64)
            return logits, hidden[last]
        gb = self.pick_graph(n)
        rows = gb or n
        self.fill_decode_layout(seqs, rows)
        src = getattr(seqs[0], "batch_src", None)
        if src is not None and all(getattr(s, "batch_src", None) is not None and s.batch_src[0] is src[0]
                                   and s.batch_src[1] == i for i, s in enumerate(seqs)):
            emb = src[0][:n].reshape(n, -1)                                   # fed by one add_requests call, in order
        else:
            emb = torch.cat([s.decode_input_embeds.reshape(1, -1) for s in seqs], 0)
        self.static_embeds[:n].copy_(emb, non_blocking=True)
        self.set_decode_context(rows)
        if gb is None:
            hidden = self.decode_body(rows)
        else:
            self.graphs[gb].replay()
            hidden = self.graph_out[gb]
        hidden = hidden[:n]
        return self.model.compute_logits(hidden), hidden

    @torch.inference_mode()
    def run(self, seqs: list[Sequence], is_prefill: bool):
        logits, hidden = self.run_model(seqs, is_prefill)
        temperatures = self.layout.d["temperatures"][:len(seqs)] if not is_prefill else self.prepare_sample(seqs)
        token_ids = self.sampler(logits, temperatures, top_k=self.top_k).tolist()              # the step's one device sync
        reset_context()
        return token_ids, hidden.clone()
