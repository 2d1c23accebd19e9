"""Code-predictor runner (reference engine/model_runner/predictor_model_runner.py:25-265): prefill
takes the 2-row ``[talker_hidden, embed(code0)]`` input of every sequence, decode step g embeds the
last sampled code with ``codec_embedding[g-1]`` and reads ``lm_head[g]``.  The embedding gather
(per-row table index) is inside the captured graph."""
from __future__ import annotations

import torch

from ...models.qwen3_tts_predictor import Qwen3TTSCodePredictorForCausalLM
from ...utils.context import reset_context
from ..sequence import PredictorSequence, Sequence  # noqa: F401  (re-exported like the reference)
from .base import ModelRunner


class PredictorModelRunner(ModelRunner):
    def pick_stack_config(self, model_config):
        self.talker_config = model_config.talker
        return model_config.predictor

    def build_model(self):
        return Qwen3TTSCodePredictorForCausalLM(self.stack_config, self.talker_config, self.rounding)

    def graph_batch_sizes(self):
        if self.config.graph_batch_sizes:
            return sorted(self.config.graph_batch_sizes)
        return list(range(1, 16)) + list(range(16, self.max_bs + 1, 16))     # predictor_model_runner.py:173

    def decode_max_context(self) -> int:
        return 32          # 2 prompt rows + 15 codes: one KV tile pair, never worth a split-KV grid

    def decode_body(self, rows: int) -> torch.Tensor:
        d = self.layout.d
        emb = self.model.get_input_embeddings(d["last_tokens"][:rows], None, d["gen_steps"][:rows])
        return self.model(emb, d["positions"][:rows])

    @torch.inference_mode()
    def run_model(self, seqs: list[Sequence], is_prefill: bool, prev_tokens: torch.Tensor | None = None):
        """``prev_tokens`` [n] (device): the codes sampled by the previous step for these rows, used instead of
        the host-side ``seq.last_token`` (deferred-token mode: the ids have not crossed PCIe yet)."""
        n = len(seqs)
        steps = [s.generation_steps for s in seqs]
        if is_prefill:
            embeds, positions = self.prepare_prefill(seqs)
            hidden = self.model(embeds, positions)
            return self.model.compute_logits(hidden, steps)
        gb = self.pick_graph(n)
        rows = gb or n
        self.fill_decode_layout(seqs, rows)
        if prev_tokens is not None:
            self.layout.d["last_tokens"][:n].copy_(prev_tokens)              # stream-ordered after the upload
        self.set_decode_context(rows)
        if gb is None:
            hidden = self.decode_body(rows)
        else:
            self.graphs[gb].replay()
            hidden = self.graph_out[gb]
        return self.model.compute_logits(hidden[:n], steps)

    @torch.inference_mode()
    def run(self, seqs: list[Sequence], is_prefill: bool) -> list[int]:
        logits = self.run_model(seqs, is_prefill)
        temperatures = self.layout.d["temperatures"][:len(seqs)] if not is_prefill else self.prepare_sample(seqs)
        token_ids = self.sampler(logits, temperatures, top_k=self.top_k).tolist()
        reset_context()
        return token_ids

    @torch.inference_mode()
    def run_deferred(self, seqs: list[Sequence], is_prefill: bool, prev_tokens: torch.Tensor | None) -> torch.Tensor:
        """``run`` without the device->host read: returns the sampled codes as a device tensor [n].  The 15-step
        predictor burst then costs ONE synchronisation (when the sequences finish) instead of one per step; the
        next step's input embedding is gathered from ``prev_tokens`` on the device."""
        logits = self.run_model(seqs, is_prefill, prev_tokens)
        temperatures = self.layout.d["temperatures"][:len(seqs)] if not is_prefill else self.prepare_sample(seqs)
        tokens = self.sampler(logits, temperatures, top_k=self.top_k)
        reset_context()
        return tokens
