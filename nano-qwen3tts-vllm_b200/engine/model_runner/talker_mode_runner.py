"""Talker runner (reference engine/model_runner/talker_mode_runner.py:23-234): decode inputs are
embeddings fed by the frame loop, ``run`` returns ``(token_ids, hidden_states)``.

Quirk kept on purpose: the decode RoPE position is ``len(seq)`` while the KV slot
index is ``len(seq) - 1`` (talker_mode_runner.py:126-128)."""
from __future__ import annotations

import numpy as np
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

    def exit(self):
        self._pf_static = None                                               # prefill graphs + their static tensors
        super().exit()

    def decode_position(self, seq: Sequence) -> int:
        return len(seq)                                                      # talker_mode_runner.py:126

    def decode_body(self, rows: int) -> torch.Tensor:
        return self.model(self.static_embeds[:rows], self.layout.d["positions"][:rows])

    # ------------------------------------------------------------------ prefill graphs
    PREFILL_GRAPH_MAX_TOKENS = 256                                            # talker_mode_runner.py:206

    def _prefill_static(self):
        """Static tensors of the single-sequence prefill graphs (talker_mode_runner.py:194-234): embeddings, one
        int64 index buffer (positions | slots) uploaded as one copy, per-length cu_seqlens."""
        st = getattr(self, "_pf_static", None)
        if st is None:
            t, dev = self.PREFILL_GRAPH_MAX_TOKENS, self.device
            st = self._pf_static = dict(
                embeds=torch.zeros(t, self.stack_config.hidden_size, dtype=BF16, device=dev),
                index=torch.zeros(2 * t, dtype=torch.int64, device=dev),
                slots=torch.zeros(t, dtype=torch.int32, device=dev),
                host=torch.zeros(2 * t, dtype=torch.int64).pin_memory(),
                graphs={})
        return st

    def _prefill_graph_ok(self, seqs: list[Sequence]) -> bool:
        return (not self.enforce_eager and len(seqs) == 1 and seqs[0].num_cached_tokens == 0
                and 1 <= len(seqs[0]) <= self.PREFILL_GRAPH_MAX_TOKENS and bool(self.graphs))

    def _run_prefill_graph(self, seq: Sequence) -> torch.Tensor:
        """Single-sequence prompt of T <= 256 tokens through a captured graph (layers + final norm, as the reference
        captures them).  The reference captures all 256 lengths at start-up; here a length is captured the first time it
        is seen (one warm run + capture) and replayed afterwards."""
        from ...utils.context import set_context
        st, t, bs = self._prefill_static(), len(seq), self.block_size
        ar = np.arange(t, dtype=np.int64)
        h = st["host"].numpy()
        h[:t] = ar
        h[self.PREFILL_GRAPH_MAX_TOKENS:self.PREFILL_GRAPH_MAX_TOKENS + t] = \
            np.asarray(seq.block_table, dtype=np.int64)[ar // bs] * bs + ar % bs
        st["index"].copy_(st["host"], non_blocking=True)
        st["slots"].copy_(st["index"][self.PREFILL_GRAPH_MAX_TOKENS:])        # int64 -> int32 on the device
        st["embeds"][:t].copy_(seq.input_embeds.reshape(-1, seq.input_embeds.shape[-1]), non_blocking=True)
        ent = st["graphs"].get(t)
        if ent is None:
            cu = torch.tensor([0, t], dtype=torch.int32, device=self.device)
            ent = st["graphs"][t] = dict(cu=cu, graph=None, out=None)
        set_context(True, ent["cu"], ent["cu"], t, t, st["slots"][:t], None, None)
        if ent["graph"] is None:
            body = lambda: self.model(st["embeds"][:t], st["index"][:t])
            body()                                                           # warm-up at this length (outside capture)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, self.graph_pool):
                ent["out"] = body()
            ent["graph"] = g
        ent["graph"].replay()
        return ent["out"]

    @torch.inference_mode()
    def run_model(self, seqs: list[Sequence], is_prefill: bool):
        n = len(seqs)
        if is_prefill:
            if self._prefill_graph_ok(seqs):
                hidden = self._run_prefill_graph(seqs[0])
            else:
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
    def run_launch(self, seqs: list[Sequence], is_prefill: bool):
        """Queue the step (layout upload, graph replay / eager pass, head, sampler, the caller's hook, the copy of the
        token ids to a pinned buffer) and return without waiting for the GPU: ``run_collect`` does the step's one wait."""
        logits, hidden = self.run_model(seqs, is_prefill)
        temperatures = self.layout.d["temperatures"][:len(seqs)] if not is_prefill else self.prepare_sample(seqs)
        tokens = self.sampler(logits, temperatures, top_k=self.top_k)
        hook = getattr(self, "post_sample_hook", None)
        if hook is not None:
            # caller work that only needs the step's DEVICE results (e.g. StepApiFrameLoop packing the predictor input and
            # starting its device->host copy) is queued before the step's one synchronisation, not after it
            hook(seqs, tokens, hidden)
        reset_context()
        n = len(seqs)
        pins = self.__dict__.setdefault("_token_pins", [None, None, 0])
        slot = pins[2] = pins[2] ^ 1
        if pins[slot] is None or pins[slot].numel() < n:
            pins[slot] = torch.empty(max(n, 64), dtype=torch.int64).pin_memory()
        host = pins[slot][:n]
        host.copy_(tokens, non_blocking=True)
        done = torch.cuda.Event()
        done.record()
        return host, done, hidden.clone()

    def run_collect(self, launched):
        host, done, hidden = launched
        done.synchronize()                                                                      # the step's one device sync
        return host.tolist(), hidden

    def run(self, seqs: list[Sequence], is_prefill: bool):
        return self.run_collect(self.run_launch(seqs, is_prefill))
