"""The per-frame talker <-> predictor loop (reference interface.py:1207-1290,
zmq/engine_loop_mp.py, workers/*), two ways:

``DeviceFrameLoop``  B200-first: talker and predictor share one process and one GPU, every stream's
    state (context lengths, block tables, last hidden state, last code) lives in HBM, and ONE CUDA
    graph holds a whole codec frame for all streams - predictor prefill + 14 predictor decode passes
    (embedding select, LM-head select and sampler inside), the 16-codebook embedding sum, the
    device-side decode layout and the talker step with its head and sampler.  One graph launch per
    frame, no host synchronisation, no per-step ``.tolist()``.

``StepApiFrameLoop``  the reference's protocol, in process: ``TalkerLLM.add_request`` /
    ``step_with_outputs`` and ``PredictorLLM.add_request`` / ``step`` with HOST tensors crossing the
    boundary both ways every frame (what the ZMQ workers exchange, protocol.py:20-29,78-91).  This is
    the end-to-end arm of bench.py.
"""
from __future__ import annotations

import torch

from .. import _native as nat
from ..layers.linear import linear
from ..sampling_params import SamplingParams
from ..utils.context import reset_context, set_context

BF16 = torch.bfloat16


class DeviceFrameLoop:
    """One codec frame per ``step()`` for ``num_streams`` streams in lock-step, all state on the device.

    Protocol of the reference's ``generate_async`` (interface.py:1207-1290), per stream: the talker's code0 ==
    codec EOS (2150) ends the stream before the frame is produced (``ignore_eos`` keeps it running, the benchmark /
    parity setting of SURVEY.md section 8d); the next talker input is the 16-codebook embedding sum plus
    ``trailing_text_hiddens[:, generation_step]`` while the trailing text lasts and ``tts_pad_embed`` after.
    Finished streams stay in the batch as padding rows (slot -1, context 0) until every stream is done.

    Test hooks: ``force(frames)`` teacher-forces the codes fed forward (the sampled ones are kept in ``sampled``),
    ``keep_logits`` copies every decision's logits into ``logit_log`` each frame."""

    def __init__(self, talker_runner, predictor_runner, num_streams: int, max_frames: int, max_prompt_len: int,
                 top_k: int = 50, talker_temperature: float = 1.0, predictor_temperature: float = 0.9,
                 scatter_pages: bool = True, seed: int = 13, ignore_eos: bool = True, codec_eos: int = 2150,
                 max_trailing: int = 0):
        self.tr, self.pr = talker_runner, predictor_runner
        self.talker, self.predictor = talker_runner.model, predictor_runner.model
        self.dev = talker_runner.device
        self.b = b = num_streams
        self.page = talker_runner.block_size
        self.top_k = top_k
        self.max_frames = max_frames
        self.eos_id = -1 if ignore_eos else int(codec_eos)
        tc, pc = talker_runner.stack_config, predictor_runner.stack_config
        self.ncb = pc.num_code_groups                                        # 16 codes per frame
        self.width = (max_prompt_len + max_frames + 1 + self.page - 1) // self.page
        need_t, need_p = b * self.width, b
        if talker_runner.kv_cache.shape[2] < need_t or predictor_runner.kv_cache.shape[2] < need_p:
            raise RuntimeError(f"KV pools too small: talker needs {need_t} pages, predictor {need_p}")
        dev = self.dev
        if scatter_pages:                                                    # non-contiguous pages, like a live pool
            perm = torch.randperm(need_t, generator=torch.Generator().manual_seed(seed)).to(torch.int32)
        else:
            perm = torch.arange(need_t, dtype=torch.int32)
        self.t_tables = perm.view(b, self.width).contiguous().to(dev)
        self.t_tables_host = perm.view(b, self.width)
        i32 = dict(dtype=torch.int32, device=dev)
        self.seq_lens = torch.zeros(b, **i32)                                # length INCLUDING the next decode token
        self.active = torch.zeros(b, **i32)
        self.gen_step = torch.zeros(b, **i32)                                # interface.py: generation_step
        self.frames_done = torch.zeros(b, **i32)
        self.extra_row = torch.zeros(b, dtype=torch.int64, device=dev)
        self.t_pos = torch.zeros(b, dtype=torch.int64, device=dev)
        self.t_slots = torch.full((b,), -1, **i32)
        self.t_ctx = torch.zeros(b, **i32)
        self.t_temp = torch.full((b,), talker_temperature, dtype=torch.float32, device=dev)
        self.p_temp = torch.full((b,), predictor_temperature, dtype=torch.float32, device=dev)
        self.embeds = torch.zeros(b, tc.hidden_size, dtype=BF16, device=dev)        # next talker input
        self.codes_t = torch.zeros(self.ncb, b, dtype=torch.int64, device=dev)      # codebook-major: row g contiguous
        self.codes = torch.zeros(b, self.ncb, dtype=torch.int64, device=dev)        # [code0 | 15 predictor codes]
        # trailing_text_hiddens[:, step] for step < trail_rows, tts_pad_embed in row trail_rows     interface.py:1282-1285
        self.trail_rows = int(max_trailing)
        self.trail = torch.zeros(b, self.trail_rows + 1, tc.hidden_size, dtype=BF16, device=dev)
        self.frames = torch.zeros(max_frames, b, self.ncb, dtype=torch.int64, device=dev)
        self.frame_idx = torch.zeros(1, dtype=torch.int64, device=dev)
        self.one = torch.ones(1, dtype=torch.int64, device=dev)
        # Exp(1) draws of the frame's 16 samplers (torch.multinomial's noise): ONE generator launch per frame
        self.vt, self.vp = tc.vocab_size, pc.vocab_size
        self.noise = torch.ones(b * (self.vt + (self.ncb - 1) * self.vp), dtype=torch.float32, device=dev)
        # predictor static layouts: stream i owns page i; prefill rows (2i, 2i+1) -> slots page*P + {0,1}
        pg = torch.arange(b, dtype=torch.int32, device=dev)
        self.p_tables = pg.view(b, 1).contiguous()
        # predictor prefill input rows (2i, 2i+1) = [last talker hidden_i, embed(code0_i)]   interface.py:1242-1252
        self.p_x = torch.zeros(2 * b, tc.hidden_size, dtype=BF16, device=dev)
        self.hidden = self.p_x.view(b, 2, -1)[:, 0]                                 # last talker hidden (strided view)
        self.p_emb0 = self.p_x.view(b, 2, -1)[:, 1]                                 # embed(code0), written by the sampler
        self.p_emb = torch.zeros(b, tc.hidden_size, dtype=BF16, device=dev)         # next predictor pass's input embedding
        self.p_pre_pos = torch.tensor([0, 1] * b, dtype=torch.int64, device=dev)
        self.p_pre_slots = (pg.view(b, 1) * self.page + torch.tensor([0, 1], dtype=torch.int32, device=dev)).view(-1).contiguous()
        self.p_cu = (torch.arange(b + 1, dtype=torch.int32, device=dev) * 2).contiguous()
        self.p_dec = []
        for g in range(1, self.ncb - 1):
            self.p_dec.append(dict(
                pos=torch.full((b,), g + 1, dtype=torch.int64, device=dev),
                ctx=torch.full((b,), g + 2, dtype=torch.int32, device=dev),
                slots=(pg * self.page + g + 1).contiguous()))
        self.graph = None
        self.launches_per_frame = None
        self.unit_noise = False
        self.frames_run = 0
        self.forced = None                 # [max_frames + 1, 16, B] codes fed forward instead of the sampled ones
        self.sampled = None                # [max_frames + 1, 16, B] what the sampler chose at each forced decision
        self.keep_logits = False
        self.logit_log: dict[int, torch.Tensor] = {}

    # ------------------------------------------------------------------ test hooks
    def force(self, frames: torch.Tensor):
        """Teacher forcing: ``frames`` [F, B, 16] are fed forward (code0 of frame f at the end of frame f - 1, the
        predictor codes inside frame f) whatever the samplers choose; the choices land in ``self.sampled``."""
        f = frames.shape[0]
        assert frames.shape[1:] == (self.b, self.ncb) and f <= self.max_frames
        self.forced = torch.zeros(self.max_frames + 1, self.ncb, self.b, dtype=torch.int64, device=self.dev)
        self.forced[:f] = frames.to(self.dev, torch.int64).permute(0, 2, 1)
        self.sampled = torch.full_like(self.forced, -1)

    def _noise_view(self, g: int) -> torch.Tensor:
        b = self.b
        if g == 0:
            return self.noise[:b * self.vt].view(b, self.vt)
        o = b * self.vt + (g - 1) * b * self.vp
        return self.noise[o:o + b * self.vp].view(b, self.vp)

    # ------------------------------------------------------------------ sampling
    def _sample(self, logits, temps, g: int):
        """Decision ``g`` of the current frame (0 = the talker's code0 of the NEXT frame): sampler.py:15-24 on the
        frame's pre-drawn Exp(1) noise; argmax(p / 1) = first maximum under ``unit_noise`` (tests).  The same launch
        gathers the embedding of the chosen code for its consumer: code0 -> row 1 of the predictor's prefill input
        (interface.py:1242), predictor code g -> the input of decode pass g (qwen3_tts_predictor.py:94-95)."""
        out = self.codes_t[g]
        if g == 0:
            table, dst = self.talker.model.codec_embedding.weight, self.p_emb0
        elif g < self.ncb - 1:
            table, dst = self.predictor.model.codec_embedding[g - 1], self.p_emb
        else:
            table, dst = None, None
        fused = table is not None and self.forced is None
        nat.topk_sample(logits, temps, self._noise_view(g), top_k=self.top_k, out=out,
                        embed_table=table if fused else None, embed_out=dst if fused else None)
        if self.keep_logits:
            buf = self.logit_log.get(g)
            if buf is None or buf.shape != logits.shape:
                buf = self.logit_log[g] = torch.empty_like(logits)
            buf.copy_(logits)
        if self.forced is not None:
            row = self.frame_idx * self.ncb + g                           # device index: graph-capturable
            self.sampled.view(-1, self.b).index_copy_(0, row, out.unsqueeze(0))
            out.copy_(self.forced.view(-1, self.b).index_select(0, row)[0])
            if table is not None:
                nat.gather_rows(table, out, out=dst)

    def _draw_noise(self):
        if not self.unit_noise:
            self.noise.exponential_(1)

    # ------------------------------------------------------------------ prompt
    @torch.inference_mode()
    def start(self, prompts: list[torch.Tensor], extra: torch.Tensor | None = None,
              trailing: list[torch.Tensor | None] | None = None):
        """Prefill every stream (packed varlen, one pass), sample the first code of each.  ``extra`` =
        tts_pad_embed [1, H] (or one row per stream); ``trailing[i]`` = trailing_text_hiddens of stream i [T_i, H]."""
        assert len(prompts) == self.b
        lens = [int(p.shape[-2]) for p in prompts]
        if max(lens) + self.max_frames + 1 > self.width * self.page:
            raise RuntimeError(f"prompt of {max(lens)} tokens + {self.max_frames} frames exceeds the {self.width}-page tables")
        self.trail.zero_()
        if extra is not None:
            self.trail.copy_(extra.to(self.dev, BF16).reshape(-1, 1, self.trail.shape[-1]).expand_as(self.trail))
        if trailing is not None:
            for i, t in enumerate(trailing):
                if t is not None and t.numel():
                    n = t.shape[-2]
                    if n > self.trail_rows:
                        raise RuntimeError(f"trailing text of {n} rows > max_trailing={self.trail_rows}")
                    self.trail[i, :n].copy_(t.reshape(n, -1))
        pos, slots, cu = [], [], [0]
        for i, n in enumerate(lens):
            pos += list(range(n))
            slots += [int(self.t_tables_host[i, t // self.page]) * self.page + t % self.page for t in range(n)]
            cu.append(cu[-1] + n)
        dev = self.dev
        cu_t = torch.tensor(cu, dtype=torch.int32, device=dev)
        set_context(True, cu_t, cu_t, max(lens), max(lens), torch.tensor(slots, dtype=torch.int32, device=dev), None, None)
        x = torch.cat([p.reshape(-1, p.shape[-1]) for p in prompts], 0).to(dev, BF16)
        h = self.talker(x, torch.tensor(pos, dtype=torch.int64, device=dev))
        last = (cu_t[1:] - 1).to(torch.int64)
        h_last = nat.gather_rows(h, last)
        self.hidden.copy_(h_last)
        logits = linear(h_last, self.talker.codec_head.weight)
        self.frame_idx.zero_()
        self._draw_noise() if not self.unit_noise else self.noise.fill_(1.0)
        self._sample(logits, self.t_temp, 0)
        reset_context()
        self.seq_lens.copy_(torch.tensor([n + 1 for n in lens], dtype=torch.int32))
        self.active.fill_(1)
        self.gen_step.zero_()
        self.frames_done.zero_()
        self.frames_run = 0
        self.first_logits = logits

    # ------------------------------------------------------------------ one frame
    def _frame(self):
        b, tr, pr = self.b, self.tr, self.pr
        talker, predictor = self.talker, self.predictor
        code0 = self.codes_t[0]
        # ---- stream state: EOS / frame budget retire a stream before its frame         interface.py:1233-1239
        nat.frame_advance(code0, self.eos_id, self.active, self.seq_lens, self.gen_step, self.frames_done,
                          self.extra_row, self.trail_rows, self.max_frames)
        self._draw_noise()
        # ---- predictor prefill on [hidden_i, embed(code0_i)] (both already in p_x)   interface.py:1242-1252
        set_context(True, self.p_cu, self.p_cu, 2, 2, self.p_pre_slots, None, None)
        qkv0 = predictor.model.layers[0].self_attn.qkv_proj.weight      # what streams after every predictor head
        h = predictor(self.p_x, self.p_pre_pos, head=0)
        last = h.view(b, 2, -1)[:, 1]
        self._sample(linear(last, predictor.lm_head[0], nxt=qkv0), self.p_temp, 1)
        # ---- 14 predictor decode passes                                 predictor_model_runner.py:134-157
        for g in range(1, self.ncb - 1):
            lay = self.p_dec[g - 1]
            set_context(False, slot_mapping=lay["slots"], context_lens=lay["ctx"], block_tables=self.p_tables,
                        max_context_len=g + 2)           # exact: no split-KV, no combine launch
            nxt = qkv0 if g + 1 < self.ncb - 1 else talker.model.layers[0].self_attn.qkv_proj.weight
            # p_emb = codec_embedding[g-1][code g], left by the sampler
            self._sample(predictor.decode_logits(self.p_emb, lay["pos"], g, nxt=nxt), self.p_temp, g + 1)
        # ---- emit the frame, build the next talker input                       interface.py:1269-1286
        self.codes.copy_(self.codes_t.t())
        self.frames.index_copy_(0, self.frame_idx, self.codes.unsqueeze(0))
        self.frame_idx.add_(self.one)
        nat.codec_embed_sum(talker.model.codec_embedding.weight, predictor.model.codec_embedding, self.codes,
                            extra=self.trail.view(-1, self.trail.shape[-1]), extra_index=self.extra_row, out=self.embeds)
        # ---- talker decode step                                        talker_mode_runner.py:116-153
        nat.decode_layout(self.seq_lens, self.t_tables, self.page, 1, 1, self.t_pos, self.t_slots, self.t_ctx)
        set_context(False, slot_mapping=self.t_slots, context_lens=self.t_ctx, block_tables=self.t_tables,
                    max_context_len=self.width * self.page)
        th = talker(self.embeds, self.t_pos)
        proj = predictor.small_to_mtp_projection
        self._sample(linear(th, talker.codec_head.weight, nxt=proj.weight if proj is not None else qkv0), self.t_temp, 0)
        self.hidden.copy_(th)
        reset_context()

    def _state(self):
        st = [self.seq_lens, self.p_x, self.p_emb, self.codes_t, self.embeds, self.frame_idx, self.frames, self.active,
              self.gen_step, self.frames_done, self.noise]
        if self.sampled is not None:
            st.append(self.sampled)
        return st

    @torch.inference_mode()
    def capture(self):
        """Capture one frame.  Device state is saved around the warm-up so capture has no side effects."""
        state = [t.clone() for t in self._state()]
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            self._frame()                                    # warm-up: lazy handles, kernel attributes
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        for t, v in zip(self._state(), state):               # the captured pass must also see a valid state
            t.copy_(v)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self._frame()
        torch.cuda.synchronize()
        for t, v in zip(self._state(), state):
            t.copy_(v)
        return self

    def step(self):
        """One codec frame for every stream: a single graph launch."""
        if self.frames_run >= self.max_frames:
            raise RuntimeError(f"DeviceFrameLoop: frame budget of {self.max_frames} frames exhausted "
                               "(frames buffer and block tables are sized for it)")
        self.frames_run += 1
        if self.graph is None:
            with torch.inference_mode():
                self._frame()
        else:
            self.graph.replay()

    def run(self, num_frames: int):
        for _ in range(num_frames):
            self.step()

    def result(self) -> torch.Tensor:
        """[frames, streams, 16] int64 on the host (synchronises).  Rows of a stream beyond its ``frame_counts()``
        entry (EOS) are padding."""
        n = int(self.frame_idx.item())
        return self.frames[:n].cpu()

    def frame_counts(self) -> list[int]:
        """Frames produced per stream (a stream stops at its codec EOS unless ``ignore_eos``)."""
        return self.frames_done.tolist()

    def result_per_stream(self) -> list[torch.Tensor]:
        fr, counts = self.result(), self.frame_counts()
        return [fr[:n, i] for i, n in enumerate(counts)]


class StepApiFrameLoop:
    """The reference's per-frame protocol (interface.py:1207-1290) over ``TalkerLLM`` / ``PredictorLLM``,
    batched over all streams, with HOST buffers at the boundary: prompts, predictor inputs and next
    talker inputs enter as pinned host tensors (host->device inside ``add_request``), hidden states
    and sampled codes come back to the host every step - the traffic the reference's ZMQ workers
    carry (protocol.py:20-29, 78-91).  The 16-codebook embedding sum runs on the GPU of the calling
    process, as in the reference (interface.py:1276-1286)."""

    def __init__(self, talker_llm, predictor_llm, tts_pad_embed: torch.Tensor, ignore_eos: bool = True,
                 codec_eos: int = 2150, batched: bool = True, device_glue: bool = True, pipelined: bool = True):
        """``batched``: feed each engine with ONE ``add_requests`` call per frame (one pinned buffer, one host->device
        copy) and read the talker's hidden states as one tensor; False = one ``add_request`` per stream, exactly the
        reference's call pattern.  Same bytes across PCIe, same results."""
        self.batched = batched
        self.talker, self.predictor = talker_llm, predictor_llm
        self.tmodel = talker_llm.model_runner.model
        self.pmodel = predictor_llm.model_runner.model
        self.dev = talker_llm.model_runner.device
        self.extra = tts_pad_embed.reshape(1, -1).to(self.dev, BF16)
        self.ignore_eos, self.codec_eos = ignore_eos, codec_eos
        self.t_sp = SamplingParams(temperature=1.0, max_tokens=1 << 30, ignore_eos=True)     # interface.py:1195
        self.p_sp = SamplingParams(temperature=0.9, max_tokens=17)                           # interface.py:1196
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        self.active: list[str] = []
        self.frames: dict[str, list[list[int]]] = {}
        self._pins: dict[str, torch.Tensor] = {}
        self.timing: dict[str, float] | None = None       # set to {} to accumulate wall time per phase (tools/e2e_profile.py)
        # Work that only needs an engine step's DEVICE results is queued from hooks the engines call before the step's
        # one synchronisation (batched mode): the predictor input [hidden, embed(code0)] is packed on the device and is
        # on its way to the pinned host buffer while the talker's token ids are being read back, and the 16-codebook sum
        # of the next talker input runs behind the predictor burst on the same stream.  The bytes still cross PCIe in
        # both directions every frame; what disappears is a second synchronisation and three launches per engine step
        # issued while the GPU idles.
        self._t_hook = None                    # (n, pinned [n, 2, H], device code0 [n]) left by the talker hook
        self._t_hook_calls = 0
        self._nxt_hook = None                  # pinned [n, H]: the next talker input, left by the predictor hook
        self._pack_dev = None
        self._fast_rids: list[str] | None = None
        self.device_glue = batched and device_glue      # False: every piece of glue starts from host values (tests)
        # pipelined: each engine's step is split into launch / collect (``step_launch`` / ``step_collect``) and the next
        # engine's work is queued - its input crossing PCIe through the pinned buffer in stream order - BEFORE the host
        # waits for the previous one, so the GPU never idles on host bookkeeping: the predictor burst of frame f is queued
        # behind the talker step of frame f (speculatively for streams whose code0 turns out to be the codec EOS: their
        # codes are dropped), and the talker step of frame f + 1 behind the burst of frame f.
        self.pipelined = self.device_glue and pipelined
        self._launched_t = None                # a talker step in flight, launched at the end of the previous frame
        if self.device_glue:
            talker_llm.model_runner.post_sample_hook = self._talker_hook
            predictor_llm.burst_hook = self._burst_hook

    def _mark(self, name: str, t0: float) -> float:
        import time
        t1 = time.perf_counter()
        if self.timing is not None:
            self.timing[name] = self.timing.get(name, 0.0) + (t1 - t0)
        return t1

    def _pinned(self, name: str, shape, dtype) -> torch.Tensor:
        """A reusable pinned host buffer (``pin_memory()`` per frame would page-lock fresh memory every time)."""
        n = 1
        for d in shape:
            n *= d
        buf = self._pins.get(name)
        if buf is None or buf.numel() < n or buf.dtype != dtype:
            with torch.inference_mode(False):      # the engine hooks run under inference_mode; these buffers are also
                buf = self._pins[name] = torch.empty(max(n, 1), dtype=dtype).pin_memory()      # written outside of it
        return buf[:n].view(*shape)

    def _to_host(self, name: str, t: torch.Tensor) -> torch.Tensor:
        out = self._pinned(name, t.shape, t.dtype)
        out.copy_(t, non_blocking=True)
        return out

    def _talker_hook(self, seqs, tokens: torch.Tensor, hidden: torch.Tensor):
        n, h = hidden.shape
        self._t_hook_calls += 1
        if self._pack_dev is None or self._pack_dev.shape[0] < n:
            self._pack_dev = torch.empty(max(n, 64), 2, h, dtype=BF16, device=self.dev)
        pk = self._pack_dev[:n]
        pk[:, 0].copy_(hidden)                                                  # interface.py:1242-1252 on the device
        nat.gather_rows(self.tmodel.model.codec_embedding.weight, tokens, out=pk[:, 1])
        host = self._pinned("pred_in", (n, 2, h), BF16)
        host.copy_(pk, non_blocking=True)
        self._t_hook = (n, host, tokens, [q.request_id for q in seqs])

    def _burst_hook(self, seqs, codes: torch.Tensor):
        self._nxt_hook = None
        fast = self._fast_rids
        if fast is None or len(seqs) != len(fast) or any(s.request_id != r for s, r in zip(seqs, fast)):
            return
        code0 = self._t_hook[2]
        codes16 = torch.cat([code0.view(-1, 1), codes], 1)                      # [n, 16] on the device
        nxt = nat.codec_embed_sum(self.tmodel.model.codec_embedding.weight, self.pmodel.model.codec_embedding, codes16,
                                  extra=self.extra)                             # interface.py:1276-1290
        self._nxt_hook = self._to_host("next_in", nxt)

    def add_streams(self, prompts: list[torch.Tensor]):
        """prompts: host bf16 [1, T, H] each.  Queues the prefill; ``step_frame`` runs it."""
        for p in prompts:
            rid = f"s{len(self.frames)}"
            self.frames[rid] = []
            self.active.append(rid)
            self.h2d_bytes += p.numel() * p.element_size()
            self.talker.add_request([p], self.t_sp, request_id=rid)

    def step_frame(self) -> int:
        """One talker step for every active stream + the predictor burst + next-input assembly.
        Returns the number of frames completed."""
        if self.pipelined:
            return self._step_frame_pipelined()
        return self._step_frame_sequential()

    def _launch_talker(self):
        self._t_hook, self._t_hook_calls, self._fast_rids, self._nxt_hook = None, 0, None, None
        return self.talker.step_launch()

    def _step_frame_pipelined(self) -> int:
        import time
        t = time.perf_counter()
        lt, self._launched_t = self._launched_t, None
        if lt is None:
            lt = self._launch_talker()
            if lt is None:
                return 0
        hooked = self._t_hook
        if hooked is None or self._t_hook_calls != 1 or hooked[0] != len(self.active) or set(hooked[3]) != set(self.active):
            # e.g. a prefill that needs several steps: this frame runs in the reference's order
            return self._step_frame_sequential(first=self.talker.step_collect(lt))
        n, pin_host, _, rids = hooked
        # ---- predictor burst of this frame, queued behind the talker step (its input: device -> pinned -> device in
        # stream order); speculative for streams about to hit the codec EOS            interface.py:1242-1268
        self.predictor.add_requests(pin_host, self.p_sp, rids)
        self._fast_rids = rids
        lp = self.predictor.step_launch()
        self.h2d_bytes += pin_host.numel() * 2
        self.d2h_bytes += pin_host.numel() * 2 + 8 * n
        t = self._mark("predictor queued behind the talker step", t)
        # ---- the talker's token ids (the wait covers the talker step only)
        outs = self.talker.step_collect(lt)[2]
        t = self._mark("talker.step_collect", t)
        code0 = [o[2][-1] for o in outs]
        assert [o[0] for o in outs] == rids
        live = [i for i, c in enumerate(code0) if self.ignore_eos or c != self.codec_eos]
        for i, rid in enumerate(rids):
            if i not in live:
                self.talker.clear_request(rid)                                  # interface.py:1233-1239
                self.predictor.clear_request(rid)
                self.active.remove(rid)
        nxt_host = self._nxt_hook
        all_live = len(live) == n and nxt_host is not None
        if all_live:
            # ---- the talker step of the NEXT frame, queued behind the burst (same stream-ordered hand-over)
            self.talker.add_requests(nxt_host.unsqueeze(1), self.t_sp, rids)
            self.h2d_bytes += nxt_host.numel() * 2
            self.d2h_bytes += nxt_host.numel() * 2
            self._launched_t = self._launch_talker()
            t = self._mark("next talker step queued behind the burst", t)
        fin, _ = self.predictor.step_collect(lp)
        done = {rid: codes for rid, _, codes in fin}
        while len(done) < n:                                                    # (a burst always finishes in one step)
            fin, _ = self.predictor.step()
            if not fin:
                break
            for rid, _, codes in fin:
                done[rid] = codes
        t = self._mark("predictor.step_collect", t)
        self.d2h_bytes += 8 * 15 * n
        rows = [[code0[i]] + done[rids[i]] for i in live]
        for row, i in zip(rows, live):
            self.frames[rids[i]].append(row)
        if live and not all_live:
            # a stream retired this frame: the survivors' next inputs are picked on the host, after the burst
            nxt_all = nxt_host if nxt_host is not None else None
            if nxt_all is None:
                codes16 = torch.tensor(rows, dtype=torch.int64)
                nxt = nat.codec_embed_sum(self.tmodel.model.codec_embedding.weight, self.pmodel.model.codec_embedding,
                                          codes16.to(self.dev, non_blocking=True), extra=self.extra)
                sel = self._to_host("next_in", nxt)
                torch.cuda.current_stream().synchronize()
                self.h2d_bytes += codes16.numel() * 8
            else:
                sel = self._pinned("next_sel", (len(live), nxt_all.shape[-1]), BF16)
                sel.copy_(nxt_all[live])
            self.h2d_bytes += sel.numel() * 2
            self.d2h_bytes += sel.numel() * 2
            self.talker.add_requests(sel.unsqueeze(1), self.t_sp, [rids[i] for i in live])
        self._mark("frames bookkeeping", t)
        return len(live)

    def _step_frame_sequential(self, first=None) -> int:
        import time
        t = time.perf_counter()
        outs = []
        if first is None:
            self._t_hook, self._t_hook_calls, self._fast_rids, self._nxt_hook = None, 0, None, None
        else:
            outs += first[2]
        while len(outs) < len(self.active):
            _, _, outs_all = self.talker.step_with_outputs()
            if not outs_all:
                break
            outs += outs_all
        if not outs:
            return 0
        t = self._mark("talker.step_with_outputs", t)
        rids = [o[0] for o in outs]
        code0 = [o[2][-1] for o in outs]
        live = [i for i, c in enumerate(code0) if self.ignore_eos or c != self.codec_eos]
        hooked = self._t_hook
        fast = hooked is not None and self._t_hook_calls == 1 and hooked[0] == len(outs) and len(live) == len(outs)
        if fast:
            # the talker hook packed [hidden, embed(code0)] on the device and its copy to the pinned buffer completed
            # with the step's synchronisation: same bytes to the host, no second round trip
            pin_host = hooked[1]
            self.d2h_bytes += pin_host.numel() * 2 + 8 * len(outs)
            self._fast_rids = rids
        else:
            last = getattr(self.talker, "last_step", None)
            whole = self.batched and last is not None and len(last[0]) == len(outs) and last[2].shape[0] == len(outs)
            hidden_dev = last[2] if whole else torch.stack([o[3] for o in outs], 0)
            hidden_host = self._to_host("hidden", hidden_dev.float())           # protocol.py:78-91 (fp32 numpy)
            self.d2h_bytes += hidden_host.numel() * 4 + 8 * len(outs)
        for i, rid in enumerate(rids):
            if i not in live:
                self.talker.clear_request(rid)                                  # interface.py:1233-1239
                self.predictor.clear_request(rid)
                self.active.remove(rid)
        if not live:
            return 0
        if not fast:
            # predictor input = [hidden, embed(code0)] per stream               interface.py:1242-1252
            c0 = torch.tensor([code0[i] for i in live], dtype=torch.int64)
            emb0 = nat.gather_rows(self.tmodel.model.codec_embedding.weight, c0.to(self.dev, non_blocking=True))
            emb0_host = self._to_host("emb0", emb0)
            torch.cuda.current_stream().synchronize()                           # hidden + embed(code0) are on the host
            pin_host = self._pinned("pred_in", (len(live), 2, hidden_host.shape[-1]), BF16)    # [n, 2, H]
            pin_host[:, 0].copy_(hidden_host if len(live) == len(outs) else hidden_host[live])
            pin_host[:, 1].copy_(emb0_host)
            self.d2h_bytes += emb0.numel() * 2
        self.h2d_bytes += pin_host.numel() * 2
        t = self._mark("hidden + embed(code0) to host, pack", t)
        if self.batched:
            self.predictor.add_requests(pin_host, self.p_sp, [rids[i] for i in live])
        else:
            for j, i in enumerate(live):
                self.predictor.add_request([pin_host[j:j + 1]], self.p_sp, request_id=rids[i])
        t = self._mark("predictor.add_requests", t)
        done: dict[str, list[int]] = {}
        while len(done) < len(live):
            fin, _ = self.predictor.step()
            for rid, _, codes in fin:
                done[rid] = codes
        t = self._mark("predictor.step (burst)", t)
        self.d2h_bytes += 8 * 15 * len(live)
        rows = [[code0[i]] + done[rids[i]] for i in live]
        nxt_host = self._nxt_hook
        if nxt_host is not None:
            # the predictor hook summed the 16 codebook embeddings behind the burst; the copy to the pinned buffer
            # completed with the burst's synchronisation
            for row, i in zip(rows, live):
                self.frames[rids[i]].append(row)
        else:
            codes16 = torch.tensor(rows, dtype=torch.int64)
            # next talker input                                                  interface.py:1276-1290
            nxt = nat.codec_embed_sum(self.tmodel.model.codec_embedding.weight, self.pmodel.model.codec_embedding,
                                      codes16.to(self.dev, non_blocking=True), extra=self.extra)
            nxt_host = self._to_host("next_in", nxt)
            for row, i in zip(rows, live):                                      # while the sum is on its way back
                self.frames[rids[i]].append(row)
            torch.cuda.current_stream().synchronize()
            self.h2d_bytes += codes16.numel() * 8
        self.h2d_bytes += nxt_host.numel() * 2
        self.d2h_bytes += nxt_host.numel() * 2
        t = self._mark("frames bookkeeping + embed sum to host", t)
        if self.batched:
            self.talker.add_requests(nxt_host.unsqueeze(1), self.t_sp, [rids[i] for i in live])
        else:
            for j, i in enumerate(live):
                self.talker.add_request([nxt_host[j:j + 1].unsqueeze(0)], self.t_sp, request_id=rids[i])
        self._mark("talker.add_requests", t)
        return len(live)

    def close(self):
        if self._launched_t is not None:                                        # a step queued ahead: let it finish
            self.talker.step_collect(self._launched_t)
            self._launched_t = None
        for rid in list(self.active):
            self.talker.clear_request(rid)
            self.predictor.clear_request(rid)
        self.active = []
        if self.device_glue:
            runner = getattr(self.talker, "model_runner", None)
            if runner is not None and getattr(runner, "post_sample_hook", None) == self._talker_hook:
                runner.post_sample_hook = None
            if getattr(self.predictor, "burst_hook", None) == self._burst_hook:
                self.predictor.burst_hook = None
