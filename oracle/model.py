"""Model-level CPU oracle: decoder layer, talker / code-predictor stacks and the
per-frame talker<->predictor loop.

TEST INFRASTRUCTURE ONLY - see ``oracle/__init__.py``.

Restates (``file:line`` under ``/root/reference/nano-qwen3tts-vllm/``):
  models/qwen3_tts_share.py:79-113,131-135,163-176   layer recipe
  models/qwen3_tts_talker.py:53-62,123-136           talker stack + codec_head
  models/qwen3_tts_predictor.py:87-124               predictor embed/head select
  layers/embed_head.py:56-61                         last-token gather in prefill
  engine/model_runner/base.py:172-234                prefill/decode batch layout
  engine/model_runner/talker_mode_runner.py:116-140  talker decode layout
  interface.py:1242-1290                             frame loop glue
Weights are a plain ``dict[str, Tensor]`` using the reference's state-dict key
names *after* ``convert_state_dict`` (prefixes stripped).
"""
from __future__ import annotations

from dataclasses import dataclass, field

import torch

from . import ops

BF16 = torch.bfloat16


@dataclass
class StackDims:
    hidden_size: int
    intermediate_size: int
    num_hidden_layers: int
    num_attention_heads: int = 16
    num_key_value_heads: int = 8
    head_dim: int = 128
    vocab_size: int = 3072
    rms_norm_eps: float = 1e-6
    rope_theta: float = 1e6
    mrope_section: list | None = None
    mrope_interleaved: bool = False
    page_size: int = 256


@dataclass
class AttnContext:
    """utils/context.py:6-15."""
    is_prefill: bool = False
    cu_seqlens_q: torch.Tensor | None = None
    cu_seqlens_k: torch.Tensor | None = None
    slot_mapping: torch.Tensor | None = None
    context_lens: torch.Tensor | None = None
    block_tables: torch.Tensor | None = None


class Stack:
    """A talker or predictor transformer stack on CPU."""

    def __init__(self, dims: StackDims, weights: dict, num_blocks: int, silu_fused_rounding=False,
                 fused_rounding=False):
        """``fused_rounding``: RMSNorm (hidden and q/k head norms) and SiLU*mul with the single store
        rounding inductor emits for the ``@torch.compile``d modules on a GPU (layernorm.py:17,29,
        qwen3_tts_share.py:116); default = the source evaluated op by op (what runs on CPU)."""
        self.d = dims
        self.w = weights
        self.fused = fused_rounding
        self.silu_fused_rounding = silu_fused_rounding or fused_rounding
        self.inv_freq = ops.make_inv_freq(dims.head_dim, dims.rope_theta)
        self.dim_modality = ops.mrope_dim_modality(dims.head_dim, dims.mrope_section, dims.mrope_interleaved)
        # base.py:158: [2, L, blocks, page, kv_heads, hd]
        self.kv_cache = torch.zeros(2, dims.num_hidden_layers, num_blocks, dims.page_size,
                                    dims.num_key_value_heads, dims.head_dim, dtype=BF16)
        self.scale = dims.head_dim ** -0.5                     # share.py:37

    # ---- share.py:79-113
    def attention(self, li, positions, x, ctx: AttnContext):
        d, w, p = self.d, self.w, f"model.layers.{li}.self_attn."
        q = ops.linear(x, w[p + "q_proj.weight"]).view(-1, d.num_attention_heads, d.head_dim)
        k = ops.linear(x, w[p + "k_proj.weight"]).view(-1, d.num_key_value_heads, d.head_dim)
        v = ops.linear(x, w[p + "v_proj.weight"]).view(-1, d.num_key_value_heads, d.head_dim)
        q, k = ops.qknorm_rope(q, k, w[p + "q_norm.weight"], w[p + "k_norm.weight"], d.rms_norm_eps,
                               positions, self.inv_freq, 1.0, self.dim_modality, self.fused)
        k_cache, v_cache = self.kv_cache[0, li], self.kv_cache[1, li]
        if ctx.slot_mapping is not None and ctx.slot_mapping.numel():
            ops.store_kvcache(k, v, k_cache, v_cache, ctx.slot_mapping)        # attention.py:76-77
        if ctx.is_prefill:
            o = ops.varlen_attn_prefill(q, k, v, ctx.cu_seqlens_q, ctx.cu_seqlens_k, self.scale,
                                        k_cache, v_cache, ctx.block_tables)
        else:
            o = ops.paged_attn_decode(q, k_cache, v_cache, ctx.context_lens, ctx.block_tables, self.scale)
        return ops.linear(o.flatten(1, -1), w[p + "o_proj.weight"])

    # ---- share.py:131-135
    def mlp(self, li, x):
        w, p = self.w, f"model.layers.{li}.mlp."
        g = ops.linear(x, w[p + "gate_proj.weight"])
        u = ops.linear(x, w[p + "up_proj.weight"])
        return ops.linear(ops.silu_and_mul(g, u, self.silu_fused_rounding), w[p + "down_proj.weight"])

    # ---- share.py:163-176
    def layer(self, li, positions, h, ctx):
        w, p = self.w, f"model.layers.{li}."
        res = h
        h = ops.rmsnorm(h, w[p + "input_layernorm.weight"], self.d.rms_norm_eps, self.fused)
        h = self.attention(li, positions, h, ctx)
        h = res + h
        res = h
        h = ops.rmsnorm(h, w[p + "post_attention_layernorm.weight"], self.d.rms_norm_eps, self.fused)
        h = self.mlp(li, h)
        return res + h

    def forward(self, embeds, positions, ctx):
        h = embeds
        for li in range(self.d.num_hidden_layers):
            h = self.layer(li, positions, h, ctx)
        return ops.rmsnorm(h, self.w["model.norm.weight"], self.d.rms_norm_eps, self.fused)


def prefill_layout(seq_lens, block_tables, page):
    """base.py:172-216 for sequences with no cached prefix."""
    positions, slots, cu = [], [], [0]
    for n, bt in zip(seq_lens, block_tables):
        positions += list(range(n))
        cu.append(cu[-1] + n)
        nblk = (n + page - 1) // page
        for i in range(nblk):
            start = bt[i] * page
            end = start + (page if i != nblk - 1 else n - (nblk - 1) * page)
            slots += list(range(start, end))
    cu_t = torch.tensor(cu, dtype=torch.int32)
    ctx = AttnContext(True, cu_t, cu_t.clone(), torch.tensor(slots, dtype=torch.int32))
    return torch.tensor(positions, dtype=torch.int64), ctx


def decode_layout(seq_lens, block_tables, page, talker: bool):
    """base.py:218-234 (predictor: position = len-1) and
    talker_mode_runner.py:116-140 (talker: position = len)."""
    positions = [n if talker else n - 1 for n in seq_lens]
    ctx_lens = list(seq_lens)
    slots = [bt[(n - 1) // page] * page + (n - 1) % page for n, bt in zip(seq_lens, block_tables)]
    width = max(len(bt) for bt in block_tables)
    bts = [list(bt) + [-1] * (width - len(bt)) for bt in block_tables]       # base.py:166-170
    ctx = AttnContext(False, None, None, torch.tensor(slots, dtype=torch.int32),
                      torch.tensor(ctx_lens, dtype=torch.int32), torch.tensor(bts, dtype=torch.int32))
    return torch.tensor(positions, dtype=torch.int64), ctx


class Talker:
    def __init__(self, dims, weights, num_blocks, **kw):
        self.stack = Stack(dims, weights, num_blocks, **kw)
        self.w = weights

    def prefill(self, embeds_list, block_tables):
        lens = [e.shape[0] for e in embeds_list]
        positions, ctx = prefill_layout(lens, block_tables, self.stack.d.page_size)
        h = self.stack.forward(torch.cat(embeds_list, 0).to(BF16), positions, ctx)
        last = ctx.cu_seqlens_q[1:].long() - 1                         # embed_head.py:57-59
        h_last = h[last].contiguous()
        return ops.linear(h_last, self.w["codec_head.weight"]), h_last

    def decode(self, embeds, seq_lens, block_tables):
        """``seq_lens`` already include the token appended after the previous step."""
        positions, ctx = decode_layout(seq_lens, block_tables, self.stack.d.page_size, talker=True)
        h = self.stack.forward(embeds.to(BF16), positions, ctx)
        return ops.linear(h, self.w["codec_head.weight"]), h


class Predictor:
    def __init__(self, dims, weights, num_blocks, talker_hidden, **kw):
        self.stack = Stack(dims, weights, num_blocks, **kw)
        self.w = weights
        self.has_proj = talker_hidden != dims.hidden_size          # predictor.py:58-61

    def _project(self, x):
        if not self.has_proj:
            return x
        return ops.linear(x, self.w["small_to_mtp_projection.weight"], self.w["small_to_mtp_projection.bias"])

    def prefill(self, embeds_list, block_tables):
        """embeds_list: per sequence [2, talker_hidden]; head 0 on the last position
        (predictor.py:110-124 with generation_steps == 0)."""
        lens = [e.shape[0] for e in embeds_list]
        positions, ctx = prefill_layout(lens, block_tables, self.stack.d.page_size)
        h = self.stack.forward(self._project(torch.cat(embeds_list, 0).to(BF16)), positions, ctx)
        last = ctx.cu_seqlens_q[1:].long() - 1
        return ops.linear(h[last], self.w["lm_head.0.weight"])

    def decode(self, last_tokens, gen_steps, seq_lens, block_tables):
        """predictor.py:94-95 (embedding g-1) and :119-121 (head g)."""
        emb = torch.stack([self.w[f"model.codec_embedding.{g - 1}.weight"][t] for t, g in zip(last_tokens, gen_steps)])
        positions, ctx = decode_layout(seq_lens, block_tables, self.stack.d.page_size, talker=False)
        h = self.stack.forward(self._project(emb.to(BF16)), positions, ctx)
        return torch.cat([ops.linear(h[i:i + 1], self.w[f"lm_head.{g}.weight"]) for i, g in enumerate(gen_steps)], 0)


def next_talker_embed(talker_w, predictor_w, code0, codes, extra):
    """interface.py:1276-1286: sum of the 16 codebook embeddings (one bf16 rounding
    from an fp32 accumulate, ``codec_hiddens.sum(1)``) plus the trailing-text /
    pad embedding (a second bf16 add)."""
    rows = [talker_w["model.codec_embedding.weight"][code0]]
    rows += [predictor_w[f"model.codec_embedding.{i}.weight"][c] for i, c in enumerate(codes)]
    s = torch.stack(rows, 0).unsqueeze(0).sum(1)                 # [1,H] bf16
    return s + extra


@dataclass
class FrameLoopResult:
    frames: list = field(default_factory=list)       # per frame: list over streams of 16 ints
    first_logits: torch.Tensor | None = None
    per_stream: list = field(default_factory=list)   # per stream: its frames (ragged once codec EOS is honoured)


def run_frames(talker: Talker, predictor: Predictor, prompts, num_frames, tts_pad_embed,
               sampler, talker_temp=1.0, predictor_temp=0.9, trailing=None, codec_eos=None):
    """Greedy/sampled frame loop for ``len(prompts)`` streams stepping in lock-step
    (interface.py:1207-1290 per stream).

    ``sampler(logits, temperatures) -> LongTensor[B]``.  Block tables are
    allocated contiguously per stream (stream s owns talker pages
    ``s*P .. s*P+P-1``); the predictor re-uses page ``s`` every frame, as the
    block manager does after ``deallocate`` (scheduler.py:73-79).
    ``trailing[s]`` = ``trailing_text_hiddens`` of stream s ``[T_s, H]`` (added at
    generation step < T_s, ``tts_pad_embed`` after, interface.py:1282-1285);
    ``codec_eos`` (2150) ends a stream before the frame whose code0 it is
    (interface.py:1233-1239); ``None`` ignores EOS.  ``res.frames`` is filled while
    every stream is alive, ``res.per_stream`` always."""
    page = talker.stack.d.page_size
    b = len(prompts)
    lens = [p.shape[0] for p in prompts]
    pages_per = (max(lens) + num_frames + 1 + page - 1) // page
    t_tables = [[s * pages_per + i for i in range(pages_per)] for s in range(b)]
    res = FrameLoopResult()
    res.per_stream = [[] for _ in range(b)]

    def table_for(n, tab):
        return tab[: (n + page - 1) // page]

    def temps(t, n):
        return torch.full((n,), t, dtype=torch.float32)

    logits, hidden = talker.prefill(prompts, [table_for(n, t) for n, t in zip(lens, t_tables)])
    res.first_logits = logits
    code0 = sampler(logits, temps(talker_temp, b)).tolist()
    lens = [n + 1 for n in lens]
    alive = list(range(b))                       # stream ids; row r of hidden / code0 belongs to alive[r]
    for step in range(num_frames):
        if codec_eos is not None:
            keep = [r for r in range(len(alive)) if code0[r] != codec_eos]
            if len(keep) != len(alive):
                hidden = hidden[keep]
                code0 = [code0[r] for r in keep]
                alive = [alive[r] for r in keep]
        if not alive:
            break
        n = len(alive)
        emb0 = talker.w["model.codec_embedding.weight"][torch.tensor(code0)]          # interface.py:1242
        pin = [torch.stack([hidden[r], emb0[r]], 0) for r in range(n)]                 # :1252
        p_tables = [[s] for s in alive]
        plogits = predictor.prefill(pin, p_tables)
        codes = [sampler(plogits, temps(predictor_temp, n)).tolist()]
        plens = [3] * n
        for g in range(1, 15):
            plogits = predictor.decode(codes[-1], [g] * n, plens, p_tables)
            codes.append(sampler(plogits, temps(predictor_temp, n)).tolist())
            plens = [m + 1 for m in plens]
        per_row = [[code0[r]] + [codes[g][r] for g in range(15)] for r in range(n)]
        for r, s in enumerate(alive):
            res.per_stream[s].append(per_row[r])
        if n == b:
            res.frames.append(per_row)
        extras = []
        for s in alive:
            t = None if trailing is None else trailing[s]
            extras.append(t[step:step + 1] if t is not None and step < t.shape[0] else tts_pad_embed)
        nxt = torch.cat([next_talker_embed(talker.w, predictor.w, per_row[r][0], per_row[r][1:], extras[r])
                         for r in range(n)], 0)
        cur = [lens[s] for s in alive]
        logits, hidden = talker.decode(nxt, cur, [table_for(lens[s], t_tables[s]) for s in alive])
        code0 = sampler(logits, temps(talker_temp, n)).tolist()
        for s in alive:
            lens[s] += 1
    return res


def synthetic_weights(talker: StackDims, predictor: StackDims, seed=1234, text_vocab=0, text_hidden=0,
                      num_code_groups=16, std=0.02):
    """Random-init bf16 state dicts: matrices / embeddings
    ``N(0, std^2)`` (``initializer_range``, config.py:299,484), norm weights
    ``1 + N(0, std^2)``, biases 0.  Keys are the reference's *converted* names
    (qwen3_tts_talker.py:81-109, qwen3_tts_predictor.py:63-85).  Drawn on the CPU
    generator one tensor at a time in a fixed order, so the same seed gives the
    same weights on any machine."""
    g = torch.Generator().manual_seed(seed)

    def mat(*shape):
        return (torch.randn(*shape, generator=g) * std).to(BF16)

    def norm(n):
        return (1.0 + torch.randn(n, generator=g) * std).to(BF16)

    def stack(d: StackDims, w: dict):
        qs, kvs = d.num_attention_heads * d.head_dim, d.num_key_value_heads * d.head_dim
        for i in range(d.num_hidden_layers):
            p = f"model.layers.{i}."
            w[p + "self_attn.q_proj.weight"] = mat(qs, d.hidden_size)
            w[p + "self_attn.k_proj.weight"] = mat(kvs, d.hidden_size)
            w[p + "self_attn.v_proj.weight"] = mat(kvs, d.hidden_size)
            w[p + "self_attn.o_proj.weight"] = mat(d.hidden_size, qs)
            w[p + "self_attn.q_norm.weight"] = norm(d.head_dim)
            w[p + "self_attn.k_norm.weight"] = norm(d.head_dim)
            w[p + "mlp.gate_proj.weight"] = mat(d.intermediate_size, d.hidden_size)
            w[p + "mlp.up_proj.weight"] = mat(d.intermediate_size, d.hidden_size)
            w[p + "mlp.down_proj.weight"] = mat(d.hidden_size, d.intermediate_size)
            w[p + "input_layernorm.weight"] = norm(d.hidden_size)
            w[p + "post_attention_layernorm.weight"] = norm(d.hidden_size)
        w["model.norm.weight"] = norm(d.hidden_size)

    tw: dict = {}
    stack(talker, tw)
    tw["model.codec_embedding.weight"] = mat(talker.vocab_size, talker.hidden_size)
    tw["codec_head.weight"] = mat(talker.vocab_size, talker.hidden_size)
    if text_vocab:
        tw["model.text_embedding.weight"] = mat(text_vocab, text_hidden)
        tw["text_projection.linear_fc1.weight"] = mat(text_hidden, text_hidden)
        tw["text_projection.linear_fc1.bias"] = torch.zeros(text_hidden, dtype=BF16)
        tw["text_projection.linear_fc2.weight"] = mat(talker.hidden_size, text_hidden)
        tw["text_projection.linear_fc2.bias"] = torch.zeros(talker.hidden_size, dtype=BF16)
    pw: dict = {}
    stack(predictor, pw)
    for j in range(num_code_groups - 1):
        pw[f"model.codec_embedding.{j}.weight"] = mat(predictor.vocab_size, talker.hidden_size)
        pw[f"lm_head.{j}.weight"] = mat(predictor.vocab_size, predictor.hidden_size)
    if talker.hidden_size != predictor.hidden_size:
        pw["small_to_mtp_projection.weight"] = mat(predictor.hidden_size, talker.hidden_size)
        pw["small_to_mtp_projection.bias"] = torch.zeros(predictor.hidden_size, dtype=BF16)
    return tw, pw
