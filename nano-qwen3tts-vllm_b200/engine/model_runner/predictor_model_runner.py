"""Code-predictor runner (reference engine/model_runner/predictor_model_runner.py:25-265): prefill
takes the 2-row ``[talker_hidden, embed(code0)]`` input of every sequence, decode step g embeds the
last sampled code with ``codec_embedding[g-1]`` and reads ``lm_head[g]``.  The embedding gather
(per-row table index) is inside the captured graph."""
from __future__ import annotations

import torch

from ... import _native as nat
from ...layers.linear import linear
from ...models.qwen3_tts_predictor import Qwen3TTSCodePredictorForCausalLM
from ...utils.context import reset_context, set_context
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

    # ------------------------------------------------------------------ the whole burst as one graph (a13)
    # A predictor request is always the same program: prefill of the 2 rows [talker hidden, embed(code0)], then one
    # decode step per remaining codebook (interface.py:1242-1268 drives it step by step).  Nothing in it depends on
    # the host, so prefill + 14 decode passes + the 15 samplers are captured ONCE per batch size and a request
    # batch that fits the pattern costs one replay and one device->host read of its [n, 15] codes.
    def capture_cudagraph(self):
        super().capture_cudagraph()
        self.capture_burst()

    def burst_codes(self) -> int:
        return self.stack_config.num_code_groups - 1

    def burst_ok(self, seqs: list[Sequence]) -> bool:
        if self.enforce_eager or not getattr(self, "burst_graphs", None) or len(seqs) > self.graph_bs[-1]:
            return False
        if self.top_k != self.burst_top_k:                                   # the sampler's k is baked into the graphs
            self.capture_burst()
        want = 2 + self.burst_codes()
        return all(s.num_tokens == 2 and s.num_cached_tokens == 0 and s.generation_steps == 0 and not s.token_ids
                   and s.max_tokens == want and len(s.block_table) == 1 for s in seqs)

    def burst_body(self, bs: int) -> torch.Tensor:
        """Static inputs: ``burst_x[:2 bs]`` (rows 2i, 2i+1 of sequence i), ``layout.d['block_tables'][:bs, 0]`` (the
        sequence's KV page, -1 = padding row), ``layout.d['temperatures'][:bs]``.  Returns codes [15, bs]."""
        model, d, page, n = self.model, self.layout.d, self.block_size, self.burst_codes()
        vocab = self.stack_config.vocab_size
        tables = d["block_tables"][:bs]
        pg = tables[:, 0]
        live = pg >= 0
        base = pg * page
        noise = self.burst_noise[:n * bs * vocab].view(n, bs, vocab)
        noise.exponential_(1)
        codes = self.burst_out[bs]
        temps = d["temperatures"][:bs]
        slots = torch.where(live.view(bs, 1), base.view(bs, 1) + self.burst_two, -1).view(-1).to(torch.int32)
        set_context(True, self.burst_cu[:bs + 1], self.burst_cu[:bs + 1], 2, 2, slots, None, None)
        qkv0 = model.model.layers[0].self_attn.qkv_proj.weight
        h = model(self.burst_x[:2 * bs], self.burst_pos2[:2 * bs], head=0)
        last = h.view(bs, 2, -1)[:, 1]
        nat.topk_sample(linear(last, model.lm_head[0], nxt=qkv0), temps, noise[0], top_k=self.top_k, out=codes[0],
                        embed_table=model.model.codec_embedding[0], embed_out=self.burst_emb[:bs])
        # the decode passes' slots / context lengths for all codebooks at once (two small launches instead of 3 per pass)
        steps = self.burst_steps[:n].view(n, 1)                              # [n, 1]: g + 1 of pass g
        slots_all = torch.where(live.view(1, bs), base.view(1, bs) + steps, -1).to(torch.int32)
        ctx_all = torch.where(live.view(1, bs), steps + 1, 0).to(torch.int32)
        for g in range(1, n):
            set_context(False, slot_mapping=slots_all[g], context_lens=ctx_all[g], block_tables=tables,
                        max_context_len=g + 2)
            last_pass = g == n - 1
            logits = model.decode_logits(self.burst_emb[:bs], self.burst_posg[g][:bs], g, nxt=None if last_pass else qkv0)
            nat.topk_sample(logits, temps, noise[g], top_k=self.top_k,
                            out=codes[g], embed_table=None if last_pass else model.model.codec_embedding[g],
                            embed_out=None if last_pass else self.burst_emb[:bs])
        reset_context()
        return codes

    @torch.inference_mode()
    def capture_burst(self):
        dev, n, mb = self.device, self.burst_codes(), self.graph_bs[-1]
        ht = self.talker_config.hidden_size
        self.burst_x = torch.zeros(2 * mb, ht, dtype=torch.bfloat16, device=dev)
        self.burst_emb = torch.zeros(mb, ht, dtype=torch.bfloat16, device=dev)
        self.burst_noise = torch.ones(n * mb * self.stack_config.vocab_size, dtype=torch.float32, device=dev)
        self.burst_two = torch.tensor([0, 1], dtype=torch.int32, device=dev).view(1, 2)
        self.burst_steps = torch.arange(1, n + 1, dtype=torch.int32, device=dev)     # row g: slot offset g + 1 of pass g
        self.burst_pos2 = torch.tensor([0, 1] * mb, dtype=torch.int64, device=dev)
        self.burst_cu = (torch.arange(mb + 1, dtype=torch.int32, device=dev) * 2).contiguous()
        self.burst_posg = [None] + [torch.full((mb,), g + 1, dtype=torch.int64, device=dev) for g in range(1, n)]
        self.burst_out = {bs: torch.zeros(n, bs, dtype=torch.int64, device=dev) for bs in self.graph_bs}
        self.burst_graphs: dict[int, torch.cuda.CUDAGraph] = {}
        self.burst_top_k = self.top_k
        h = self.layout.begin()
        h["block_tables"][:] = -1
        self.layout.upload()
        for bs in reversed(self.graph_bs):
            self.burst_body(bs)                                             # warm-up (packs lm_head images, lazy handles)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, self.graph_pool):
                self.burst_body(bs)
            self.burst_graphs[bs] = g
        torch.cuda.synchronize()

    @torch.inference_mode()
    def run_burst(self, seqs: list[Sequence]) -> torch.Tensor:
        """All codes of ``seqs`` (fresh 2-row requests, see ``burst_ok``): device tensor [n, 15]."""
        n = len(seqs)
        bs = self.pick_graph(n)
        h = self.layout.begin()
        h["block_tables"][:n, 0] = [s.block_table[0] for s in seqs]
        h["block_tables"][n:bs, 0] = -1
        h["temperatures"][:n] = [s.temperature for s in seqs]
        h["temperatures"][n:bs] = 1.0
        self.layout.upload()
        src = getattr(seqs[0], "batch_src", None)
        if src is not None and all(getattr(s, "batch_src", None) is not None and s.batch_src[0] is src[0]
                                   and s.batch_src[1] == i for i, s in enumerate(seqs)):
            self.burst_x[:2 * n].copy_(src[0][:n].reshape(2 * n, -1))        # one batched upload: already contiguous
        else:
            self.burst_x[:2 * n].copy_(torch.cat([s.input_embeds.reshape(2, -1) for s in seqs], 0))
        self.burst_graphs[bs].replay()
        return self.burst_out[bs][:, :n].t()

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
