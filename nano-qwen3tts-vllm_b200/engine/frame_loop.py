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
    def __init__(self, talker_runner, predictor_runner, num_streams: int, max_frames: int, max_prompt_len: int,
                 top_k: int = 50, talker_temperature: float = 1.0, predictor_temperature: float = 0.9,
                 scatter_pages: bool = True, seed: int = 13):
        self.tr, self.pr = talker_runner, predictor_runner
        self.talker, self.predictor = talker_runner.model, predictor_runner.model
        self.dev = talker_runner.device
        self.b = b = num_streams
        self.page = talker_runner.block_size
        self.top_k = top_k
        self.max_frames = max_frames
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
        self.seq_lens = torch.zeros(b, dtype=torch.int32, device=dev)       # length INCLUDING the next decode token
        self.t_pos = torch.zeros(b, dtype=torch.int64, device=dev)
        self.t_slots = torch.full((b,), -1, dtype=torch.int32, device=dev)
        self.t_ctx = torch.zeros(b, dtype=torch.int32, device=dev)
        self.t_temp = torch.full((b,), talker_temperature, dtype=torch.float32, device=dev)
        self.p_temp = torch.full((b,), predictor_temperature, dtype=torch.float32, device=dev)
        self.embeds = torch.zeros(b, tc.hidden_size, dtype=BF16, device=dev)        # next talker input
        self.hidden = torch.zeros(b, tc.hidden_size, dtype=BF16, device=dev)        # last talker hidden
        self.codes_t = torch.zeros(self.ncb, b, dtype=torch.int64, device=dev)      # codebook-major: row g contiguous
        self.codes = torch.zeros(b, self.ncb, dtype=torch.int64, device=dev)        # [code0 | 15 predictor codes]
        self.extra = torch.zeros(1, tc.hidden_size, dtype=BF16, device=dev)         # tts_pad_embed / trailing text
        self.frames = torch.zeros(max_frames, b, self.ncb, dtype=torch.int64, device=dev)
        self.frame_idx = torch.zeros(1, dtype=torch.int64, device=dev)
        self.one = torch.ones(1, dtype=torch.int64, device=dev)
        # predictor static layouts: stream i owns page i; prefill rows (2i, 2i+1) -> slots page*P + {0,1}
        pg = torch.arange(b, dtype=torch.int32, device=dev)
        self.p_tables = pg.view(b, 1).contiguous()
        self.p_x = torch.zeros(2 * b, tc.hidden_size, dtype=BF16, device=dev)
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

    # ------------------------------------------------------------------ sampling
    def _sample(self, logits, temps, out):
        if self.unit_noise:          # tests: argmax(p / 1) = first maximum, so exact bf16 ties do not depend on the RNG
            noise = torch.ones(logits.shape, dtype=torch.float32, device=self.dev)
        else:
            noise = torch.empty(logits.shape, dtype=torch.float32, device=self.dev).exponential_(1)
        nat.topk_sample(logits, temps, noise, top_k=self.top_k, out=out)

    # ------------------------------------------------------------------ prompt
    @torch.inference_mode()
    def start(self, prompts: list[torch.Tensor], extra: torch.Tensor | None = None):
        """Prefill every stream (packed varlen, one pass), sample the first code of each."""
        assert len(prompts) == self.b
        lens = [int(p.shape[-2]) for p in prompts]
        if extra is not None:
            self.extra.copy_(extra.reshape(1, -1))
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
        self.hidden.copy_(nat.gather_rows(h, last))
        logits = linear(self.hidden, self.talker.codec_head.weight)
        self._sample(logits, self.t_temp, self.codes_t[0])
        reset_context()
        self.seq_lens.copy_(torch.tensor([n + 1 for n in lens], dtype=torch.int32))
        self.frame_idx.zero_()
        self.first_logits = logits

    # ------------------------------------------------------------------ one frame
    def _frame(self):
        b, tr, pr = self.b, self.tr, self.pr
        talker, predictor = self.talker, self.predictor
        code0 = self.codes_t[0]
        # ---- predictor prefill on [hidden_i, embed(code0_i)]                   interface.py:1242-1252
        x3 = self.p_x.view(b, 2, -1)
        x3[:, 0].copy_(self.hidden)
        nat.gather_rows(talker.model.codec_embedding.weight, code0, out=x3[:, 1])
        set_context(True, self.p_cu, self.p_cu, 2, 2, self.p_pre_slots, None, None)
        h = predictor(self.p_x, self.p_pre_pos)
        last = h.view(b, 2, -1)[:, 1]
        self._sample(linear(last, predictor.lm_head[0]), self.p_temp, self.codes_t[1])
        # ---- 14 predictor decode passes                                 predictor_model_runner.py:134-157
        for g in range(1, self.ncb - 1):
            lay = self.p_dec[g - 1]
            emb = nat.gather_rows(predictor.model.codec_embedding[g - 1], self.codes_t[g])
            set_context(False, slot_mapping=lay["slots"], context_lens=lay["ctx"], block_tables=self.p_tables,
                        max_context_len=g + 2)           # exact: no split-KV, no combine launch
            h = predictor(emb, lay["pos"])
            self._sample(linear(h, predictor.lm_head[g]), self.p_temp, self.codes_t[g + 1])
        # ---- emit the frame, build the next talker input                       interface.py:1269-1286
        self.codes.copy_(self.codes_t.t())
        self.frames.index_copy_(0, self.frame_idx, self.codes.unsqueeze(0))
        self.frame_idx.add_(self.one)
        nat.codec_embed_sum(talker.model.codec_embedding.weight, predictor.model.codec_embedding, self.codes,
                            extra=self.extra, out=self.embeds)
        # ---- talker decode step                                        talker_mode_runner.py:116-153
        nat.decode_layout(self.seq_lens, self.t_tables, self.page, 1, 1, self.t_pos, self.t_slots, self.t_ctx)
        set_context(False, slot_mapping=self.t_slots, context_lens=self.t_ctx, block_tables=self.t_tables,
                    max_context_len=self.width * self.page)
        th = talker(self.embeds, self.t_pos)
        self.hidden.copy_(th)
        self._sample(linear(th, talker.codec_head.weight), self.t_temp, self.codes_t[0])
        reset_context()

    @torch.inference_mode()
    def capture(self):
        """Capture one frame.  Device state is saved around the warm-up so capture has no side effects."""
        state = [t.clone() for t in (self.seq_lens, self.hidden, self.codes_t, self.embeds, self.frame_idx, self.frames)]
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            self._frame()                                    # warm-up: lazy handles, kernel attributes
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self._frame()
        torch.cuda.synchronize()
        for t, v in zip((self.seq_lens, self.hidden, self.codes_t, self.embeds, self.frame_idx, self.frames), state):
            t.copy_(v)
        return self

    def step(self):
        """One codec frame for every stream: a single graph launch."""
        if self.graph is None:
            with torch.inference_mode():
                self._frame()
        else:
            self.graph.replay()

    def run(self, num_frames: int):
        for _ in range(num_frames):
            self.step()

    def result(self) -> torch.Tensor:
        """[frames, streams, 16] int64 on the host (synchronises)."""
        n = int(self.frame_idx.item())
        return self.frames[:n].cpu()


class StepApiFrameLoop:
    """The reference's per-frame protocol (interface.py:1207-1290) over ``TalkerLLM`` / ``PredictorLLM``,
    batched over all streams, with HOST buffers at the boundary: prompts, predictor inputs and next
    talker inputs enter as pinned host tensors (host->device inside ``add_request``), hidden states
    and sampled codes come back to the host every step - the traffic the reference's ZMQ workers
    carry (protocol.py:20-29, 78-91).  The 16-codebook embedding sum runs on the GPU of the calling
    process, as in the reference (interface.py:1276-1286)."""

    def __init__(self, talker_llm, predictor_llm, tts_pad_embed: torch.Tensor, ignore_eos: bool = True,
                 codec_eos: int = 2150):
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
        outs = []
        while len(outs) < len(self.active):
            _, _, outs_all = self.talker.step_with_outputs()
            if not outs_all:
                break
            outs += outs_all
        if not outs:
            return 0
        rids = [o[0] for o in outs]
        code0 = [o[2][-1] for o in outs]
        hidden_host = torch.stack([o[3] for o in outs], 0).float().cpu()       # protocol.py:78-91 (fp32 numpy)
        self.d2h_bytes += hidden_host.numel() * 4 + 8 * len(outs)
        live = [i for i, c in enumerate(code0) if self.ignore_eos or c != self.codec_eos]
        for i, rid in enumerate(rids):
            if i not in live:
                self.talker.clear_request(rid)                                  # interface.py:1233-1239
                self.predictor.clear_request(rid)
                self.active.remove(rid)
        if not live:
            return 0
        # predictor input = [hidden, embed(code0)] per stream                   interface.py:1242-1252
        c0 = torch.tensor([code0[i] for i in live], dtype=torch.int64)
        emb0 = nat.gather_rows(self.tmodel.model.codec_embedding.weight, c0.to(self.dev))
        pin_host = torch.stack([hidden_host[live].to(BF16), emb0.cpu()], 1).pin_memory()       # [n, 2, H]
        self.d2h_bytes += emb0.numel() * 2
        for j, i in enumerate(live):
            self.h2d_bytes += pin_host[j].numel() * 2
            self.predictor.add_request([pin_host[j:j + 1]], self.p_sp, request_id=rids[i])
        done: dict[str, list[int]] = {}
        while len(done) < len(live):
            fin, _ = self.predictor.step()
            for rid, _, codes in fin:
                done[rid] = codes
        self.d2h_bytes += 8 * 15 * len(live)
        codes16 = torch.tensor([[code0[i]] + done[rids[i]] for i in live], dtype=torch.int64)
        for row, i in zip(codes16.tolist(), live):
            self.frames[rids[i]].append(row)
        # next talker input                                                      interface.py:1276-1290
        nxt = nat.codec_embed_sum(self.tmodel.model.codec_embedding.weight, self.pmodel.model.codec_embedding,
                                  codes16.to(self.dev), extra=self.extra)
        nxt_host = nxt.cpu().pin_memory()
        self.h2d_bytes += codes16.numel() * 8 + nxt_host.numel() * 2
        self.d2h_bytes += nxt_host.numel() * 2
        for j, i in enumerate(live):
            self.talker.add_request([nxt_host[j:j + 1].unsqueeze(0)], self.t_sp, request_id=rids[i])
        return len(live)

    def close(self):
        for rid in list(self.active):
            self.talker.clear_request(rid)
            self.predictor.clear_request(rid)
        self.active = []
