"""Pin the CPU oracle against the LIVE reference modules and write golden vectors.

TEST INFRASTRUCTURE ONLY - see ``oracle/__init__.py``.

Runs only in the build container (needs ``/root/reference``); the GPU box never
executes this file.  It imports the reference package through a symlink
(``/tmp/refpkg/nano_qwen3tts_vllm`` -> ``/root/reference/nano-qwen3tts-vllm``),
with ``TORCHDYNAMO_DISABLE=1`` (inductor's CPU build is broken in this image) and
a world-size-1 gloo group (``layers/linear.py:23-24`` calls ``dist.get_rank()``),
feeds seeded inputs to the reference's own modules, asserts that ``oracle.ops`` /
``oracle.model`` reproduce every output BIT-EXACTLY, and stores inputs + reference
outputs under ``tests/golden/*.pt`` for the CPU and GPU test-suites.

Two reference ops are CUDA-only (Triton ``store_kvcache``, flash-attn): for the
model-level vectors ``Attention.forward`` is replaced by the oracle restatement of
``layers/attention.py:73-96`` - those two ops stay "parity unpinned" here and are
cross-checked against the flash-attn wheel on the GPU box instead.

    TORCHDYNAMO_DISABLE=1 python oracle/pin_against_reference.py
"""
from __future__ import annotations

import os
import sys
import types

os.environ.setdefault("TORCHDYNAMO_DISABLE", "1")

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")
REF_SRC = "/root/reference/nano-qwen3tts-vllm"
BF16 = torch.bfloat16


def import_reference():
    link_dir = "/tmp/refpkg"
    os.makedirs(link_dir, exist_ok=True)
    link = os.path.join(link_dir, "nano_qwen3tts_vllm")
    if not os.path.islink(link):
        os.symlink(REF_SRC, link)
    sys.path.insert(0, link_dir)
    if not dist.is_initialized():
        dist.init_process_group("gloo", init_method="tcp://127.0.0.1:29731", world_size=1, rank=0)


def eq(a, b, what):
    assert a.dtype == b.dtype and a.shape == b.shape, (what, a.dtype, b.dtype, a.shape, b.shape)
    if not torch.equal(a, b):
        d = (a.float() - b.float()).abs().max().item()
        raise AssertionError(f"oracle != reference for {what}: max abs diff {d}")


def randn(gen, *shape, std=1.0, dtype=BF16):
    return (torch.randn(*shape, generator=gen) * std).to(dtype)


def pin_ops():
    from oracle import ops
    from nano_qwen3tts_vllm.layers.layernorm import Qwen3TTSRMSNorm
    from nano_qwen3tts_vllm.layers.rotary_embedding import get_rope
    from nano_qwen3tts_vllm.layers.activation import Silu
    from nano_qwen3tts_vllm.layers.sampler import Sampler

    g = torch.Generator().manual_seed(20261104)
    out = {}

    # ---- RMSNorm (hidden and per-head shapes) -------------- layernorm.py:18-41
    for name, shape in (("rms_hidden", (7, 2048)), ("rms_head", (5, 16, 128)), ("rms_1024", (3, 1024))):
        x = randn(g, *shape, std=1.5)
        res = randn(g, *shape, std=0.7)
        m = Qwen3TTSRMSNorm(shape[-1], eps=1e-6).to(BF16)
        m.weight.data = (1 + 0.05 * torch.randn(shape[-1], generator=g)).to(BF16)
        y = m(x)
        y2, r2 = m(x, res)
        eq(ops.rmsnorm(x, m.weight.data, 1e-6), y, name)
        oy2, or2 = ops.add_rmsnorm(x, res, m.weight.data, 1e-6)
        eq(oy2, y2, name + ".add"); eq(or2, r2, name + ".add.res")
        out[name] = dict(x=x, res=res, w=m.weight.data.clone(), eps=1e-6, y=y, y_add=y2, res_add=r2)

    # ---- RoPE ------------------------------------------ rotary_embedding.py:101-237
    t = 9
    q = randn(g, t, 16, 128)
    k = randn(g, t, 8, 128)
    pos = torch.tensor([0, 1, 2, 63, 64, 577, 1023, 4095, 4351], dtype=torch.int64)
    rope1 = get_rope(128, 128, 32768, 1e6, None)
    rq, rk = rope1(pos, q, k)
    oq, ok = ops.rope(pos, q, k, rope1.inv_freq)
    eq(oq, rq, "rope1d.q"); eq(ok, rk, "rope1d.k")
    out["rope_1d"] = dict(q=q, k=k, positions=pos, theta=1e6, inv_freq=rope1.inv_freq.clone(), q_out=rq, k_out=rk)
    for inter in (True, False):
        sc = {"mrope_section": [24, 20, 20], "interleaved": inter}
        ropem = get_rope(128, 128, 32768, 1e6, sc)
        modality = ops.mrope_dim_modality(128, sc["mrope_section"], inter)
        # (a) what the runner really feeds: 1-D positions (talker_mode_runner.py:134)
        rq, rk = ropem(pos, q, k)
        oq, ok = ops.rope(pos, q, k, ropem.inv_freq, 1.0, modality)
        eq(oq, rq, f"mrope{inter}.1d.q"); eq(ok, rk, f"mrope{inter}.1d.k")
        eq(oq, out["rope_1d"]["q_out"], "mrope collapses to 1-D rope")
        # (b) three distinct rows [3, 1, T]
        pos3 = torch.stack([pos, pos // 2 + 3, pos // 3 + 11])
        rq3, rk3 = ropem(pos3.unsqueeze(1), q, k)
        oq3, ok3 = ops.rope(pos3, q, k, ropem.inv_freq, 1.0, modality)
        eq(oq3, rq3, f"mrope{inter}.3d.q"); eq(ok3, rk3, f"mrope{inter}.3d.k")
        out[f"mrope_{'interleaved' if inter else 'sectioned'}"] = dict(
            q=q, k=k, positions3=pos3, mrope_section=[24, 20, 20], interleaved=inter,
            modality=torch.tensor(modality, dtype=torch.int32), inv_freq=ropem.inv_freq.clone(),
            q_out=rq3, k_out=rk3)

    # ---- SiLU / gated product --------------- activation.py:12-13, share.py:131-135
    gate = randn(g, 6, 3072, std=2.0)
    up = randn(g, 6, 3072, std=1.0)
    s = Silu()(gate)
    eq(ops.silu(gate), s, "silu")
    eq(ops.silu_and_mul(gate, up), s * up, "silu_and_mul eager")
    out["silu"] = dict(gate=gate, up=up, silu=s, silu_mul_eager=s * up,
                       silu_mul_fused=ops.silu_and_mul(gate, up, fused_rounding=True))

    # ---- Sampler ------------------------------------------------ sampler.py:15-24
    sampler = Sampler()
    for name, (b, v, temp, topk) in dict(talker=(6, 3072, 1.0, 50), predictor=(5, 2048, 0.9, 50),
                                         greedy=(4, 3072, 1.0, 1)).items():
        logits = randn(g, b, v, std=3.0)
        logits[0, 100:140] = logits[0, 7]          # ties straddling the k-th value are kept
        temps = torch.full((b,), temp, dtype=torch.float32)
        temps[-1] = 0.5
        torch.manual_seed(4242)
        ref_idx = sampler(logits, temps, top_k=topk)
        torch.manual_seed(4242)
        noise = torch.empty(b, v, dtype=torch.float32).exponential_(1)   # what multinomial draws
        probs = ops.topk_probs(logits, temps, topk)
        idx = ops.sample_from_probs(probs, noise)
        eq(idx, ref_idx, "sampler." + name)
        out["sampler_" + name] = dict(logits=logits, temperatures=temps, noise=noise, top_k=topk,
                                      probs=probs, index=ref_idx)
    return out


# ------------------------------------------------------------------ model level
def small_dims():
    from oracle.model import StackDims
    talker = StackDims(hidden_size=256, intermediate_size=512, num_hidden_layers=2, vocab_size=3072,
                       rope_theta=1e6, mrope_section=[24, 20, 20], mrope_interleaved=True, page_size=256)
    predictor = StackDims(hidden_size=128, intermediate_size=384, num_hidden_layers=2, vocab_size=2048,
                          rope_theta=1e6, page_size=256)
    return talker, predictor


def ns_config(d, **extra):
    rope_scaling = None
    if d.mrope_section is not None:
        rope_scaling = {"mrope_section": d.mrope_section, "interleaved": d.mrope_interleaved}
    return types.SimpleNamespace(
        vocab_size=d.vocab_size, hidden_size=d.hidden_size, intermediate_size=d.intermediate_size,
        num_hidden_layers=d.num_hidden_layers, num_attention_heads=d.num_attention_heads,
        num_key_value_heads=d.num_key_value_heads, head_dim=d.head_dim, rms_norm_eps=d.rms_norm_eps,
        rope_theta=d.rope_theta, rope_scaling=rope_scaling, attention_bias=False,
        max_position_embeddings=32768, hidden_act="silu", **extra)


def patch_reference_attention(stack_of):
    """Replace the CUDA-only body of Attention.forward (attention.py:73-96) by the
    oracle restatement, keeping the reference's own Context plumbing."""
    from oracle import ops
    from nano_qwen3tts_vllm.layers import attention as ref_attn
    from nano_qwen3tts_vllm.utils.context import get_context

    def forward(self, q, k, v):
        ctx = get_context()
        if self.k_cache.numel() and self.v_cache.numel():
            ops.store_kvcache(k, v, self.k_cache, self.v_cache, ctx.slot_mapping)
        if ctx.is_prefill:
            return ops.varlen_attn_prefill(q, k, v, ctx.cu_seqlens_q, ctx.cu_seqlens_k, self.scale,
                                           self.k_cache, self.v_cache, ctx.block_tables)
        return ops.paged_attn_decode(q, self.k_cache, self.v_cache, ctx.context_lens, ctx.block_tables, self.scale)

    ref_attn.Attention.forward = forward


def bind_kv(model, kv_cache):
    """engine/model_runner/base.py:158-164."""
    layer_id = 0
    for module in model.modules():
        if hasattr(module, "k_cache") and hasattr(module, "v_cache"):
            module.k_cache = kv_cache[0, layer_id]
            module.v_cache = kv_cache[1, layer_id]
            layer_id += 1


def pin_models():
    from oracle import model as om
    from nano_qwen3tts_vllm.models.qwen3_tts_talker import Qwen3TTSTalkerForCausalLM
    from nano_qwen3tts_vllm.models.qwen3_tts_predictor import Qwen3TTSCodePredictorForCausalLM
    from nano_qwen3tts_vllm.utils.context import set_context, reset_context

    td, pd = small_dims()
    seed = 1234
    tw, pw = om.synthetic_weights(td, pd, seed=seed, text_vocab=64, text_hidden=96)
    torch.set_default_dtype(BF16)                       # base.py:55
    tcfg = ns_config(td, text_vocab_size=64, text_hidden_size=96, num_code_groups=16)
    pcfg = ns_config(pd, num_code_groups=16, talker_hidden_size=td.hidden_size)
    ref_t = Qwen3TTSTalkerForCausalLM(tcfg)
    ref_p = Qwen3TTSCodePredictorForCausalLM(pcfg, tcfg)
    torch.set_default_dtype(torch.float32)
    full_sd = {"talker." + k: v for k, v in tw.items()}
    full_sd.update({"talker.code_predictor." + k: v for k, v in pw.items()})
    ref_t.load_state_dict(full_sd)
    ref_p.load_state_dict(full_sd)
    patch_reference_attention(None)

    nblk = 8
    o_t = om.Talker(td, tw, nblk)
    o_p = om.Predictor(pd, pw, nblk, td.hidden_size)
    bind_kv(ref_t, torch.zeros_like(o_t.stack.kv_cache))
    bind_kv(ref_p, torch.zeros_like(o_p.stack.kv_cache))

    g = torch.Generator().manual_seed(777)
    out = dict(seed=seed, num_blocks=nblk)
    with torch.inference_mode():
        # ---- talker prefill: two ragged sequences, non-contiguous pages
        lens = [11, 6]
        tables = [[5], [2]]
        prompts = [randn(g, n, td.hidden_size, std=0.05) for n in lens]
        pos, ctx = om.prefill_layout(lens, tables, 256)
        set_context(True, ctx.cu_seqlens_q, ctx.cu_seqlens_k, max(lens), max(lens), ctx.slot_mapping, None, None)
        h = ref_t(torch.cat(prompts, 0), pos)
        logits = ref_t.compute_logits(h)
        hid = h[ctx.cu_seqlens_q[1:].long() - 1]
        reset_context()
        ol, oh = o_t.prefill(prompts, tables)
        eq(ol, logits, "talker.prefill.logits"); eq(oh, hid, "talker.prefill.hidden")
        out["talker_prefill"] = dict(prompts=prompts, tables=tables, logits=logits, hidden=hid)
        # ---- talker decode: 3 steps (position = len, slot = len-1: talker_mode_runner.py:126-128)
        steps = []
        cur = list(lens)
        for _ in range(3):
            cur = [n + 1 for n in cur]
            emb = randn(g, 2, td.hidden_size, std=0.05)
            pos, ctx = om.decode_layout(cur, tables, 256, talker=True)
            set_context(False, slot_mapping=ctx.slot_mapping, context_lens=ctx.context_lens, block_tables=ctx.block_tables)
            h = ref_t(emb, pos)
            logits = ref_t.compute_logits(h)
            reset_context()
            ol, oh = o_t.decode(emb, cur, tables)
            eq(ol, logits, "talker.decode.logits"); eq(oh, h, "talker.decode.hidden")
            steps.append(dict(embeds=emb, seq_lens=list(cur), logits=logits, hidden=h))
        out["talker_decode"] = steps
        eq(o_t.stack.kv_cache, torch.stack([torch.stack([m.self_attn.attn.k_cache for m in ref_t.model.layers]),
                                            torch.stack([m.self_attn.attn.v_cache for m in ref_t.model.layers])]),
           "talker.kv_cache")
        out["talker_kv_pages"] = o_t.stack.kv_cache[:, :, [5, 2], :16].clone()

        # ---- predictor prefill (2 tokens / seq) and decode with per-seq generation_steps
        b = 3
        ptables = [[1], [4], [0]]
        pin = [randn(g, 2, td.hidden_size, std=0.5) for _ in range(b)]
        pos, ctx = om.prefill_layout([2] * b, ptables, 256)
        set_context(True, ctx.cu_seqlens_q, ctx.cu_seqlens_k, 2, 2, ctx.slot_mapping, None, None)
        emb = ref_p.get_input_embeddings(None, torch.cat(pin, 0).unsqueeze(0), [0] * b)   # predictor.py:89-91
        h = ref_p(emb.view(-1, emb.size(-1)), pos)
        logits = ref_p.compute_logits(h, [0] * b)
        reset_context()
        ol = o_p.prefill(pin, ptables)
        eq(ol, logits, "predictor.prefill.logits")
        out["predictor_prefill"] = dict(inputs=pin, tables=ptables, logits=logits)
        steps = []
        plens = [2] * b
        for gstep in (1, 2, 14):
            plens = [n + 1 for n in plens]
            toks = torch.randint(0, pd.vocab_size, (b,), generator=g).tolist()
            gsteps = [gstep] * b
            pos, ctx = om.decode_layout(plens, ptables, 256, talker=False)
            set_context(False, slot_mapping=ctx.slot_mapping, context_lens=ctx.context_lens, block_tables=ctx.block_tables)
            emb = ref_p.get_input_embeddings(torch.tensor(toks), None, gsteps)
            h = ref_p(emb, pos)
            logits = ref_p.compute_logits(h, gsteps)
            reset_context()
            ol = o_p.decode(toks, gsteps, plens, ptables)
            eq(ol, logits, f"predictor.decode.g{gstep}")
            steps.append(dict(tokens=toks, gen_steps=gsteps, seq_lens=list(plens), logits=logits))
        out["predictor_decode"] = steps
    return out


# ------------------------------------------------------------------ a11: KV-pool memory policy
def pin_memory_policy():
    """Run the reference's own memory-policy code on tables of inputs and assert that
    nano_qwen3tts_vllm_b200.engine.memory reproduces every number exactly.

    ``interface.py`` cannot be imported (qwen_tts / librosa / soundfile are absent), so the three pure
    functions ``_estimate_model_params`` / ``_kv_block_bytes`` / ``_compute_memory_split``
    (interface.py:57-197) are lifted from its source by ``ast`` and executed unmodified with a stand-in for
    ``torch.cuda.get_device_properties``; ``ModelRunner.allocate_kv_cache`` (engine/model_runner/base.py:131-164)
    is the reference's unbound method called on a stub runner whose ``torch`` exposes the allocator numbers of
    the table row."""
    import ast
    import json
    import logging
    import tempfile

    import nano_qwen3tts_vllm_b200.engine.memory as mem
    from nano_qwen3tts_vllm_b200.config import ModelConfig, qwen3_tts_config
    from nano_qwen3tts_vllm.config import Qwen3TTSConfig
    import nano_qwen3tts_vllm.engine.model_runner.base as ref_base

    src = open(os.path.join(REF_SRC, "interface.py")).read()
    wanted = {"_estimate_model_params", "_kv_block_bytes", "_compute_memory_split"}
    tree = ast.parse(src)
    body = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in wanted]
    assert {n.name for n in body} == wanted
    vram = {"total": 0}
    fake_torch = types.SimpleNamespace(cuda=types.SimpleNamespace(
        get_device_properties=lambda i: types.SimpleNamespace(total_memory=vram["total"])))
    ns = {"os": os, "json": json, "torch": fake_torch, "Qwen3TTSConfig": Qwen3TTSConfig,
          "logger": logging.getLogger("ref.interface")}
    exec(compile(ast.Module(body=body, type_ignores=[]), "interface.py[memory policy]", "exec"), ns)

    rows = {"split": [], "blocks": []}
    for size in ("1.7B", "0.6B"):
        mc = qwen3_tts_config(size)
        raw = mc.to_dict()
        for stack in (raw["talker_config"], raw["talker_config"]["code_predictor_config"]):
            stack.pop("text_vocab_size", None) if not stack.get("text_vocab_size") else None
            stack.pop("text_hidden_size", None) if not stack.get("text_hidden_size") else None
        with tempfile.TemporaryDirectory() as d:
            with open(os.path.join(d, "config.json"), "w") as f:
                json.dump(raw, f)
            ref_cfg = Qwen3TTSConfig(**raw)
            for cfg_ref, cfg_ours in ((ref_cfg.talker_config, mc.talker),
                                      (ref_cfg.talker_config.code_predictor_config, mc.predictor)):
                assert ns["_estimate_model_params"](cfg_ref) == mem.estimate_model_params(cfg_ours)
                assert ns["_kv_block_bytes"](cfg_ref) == mem.kv_block_bytes(cfg_ours)
            for total in (24 << 30, 80 << 30, 85_899_345_920, 191_514_542_080):
                for util in (0.02, 0.05, 0.3, 0.5, 0.9, 0.95):
                    vram["total"] = total
                    want = ns["_compute_memory_split"](d, util)
                    got = mem.compute_memory_split(mc, util, total)
                    for k in ("talker_util", "pred_util", "process_gpu_memory_fraction"):
                        assert got[k] == want[k], (size, total, util, k, got[k], want[k])
                    rows["split"].append(dict(model=size, total_vram=total, gpu_memory_utilization=util,
                                              talker_util=want["talker_util"], pred_util=want["pred_util"],
                                              process_gpu_memory_fraction=want["process_gpu_memory_fraction"]))

    # ---- allocate_kv_cache: the reference method on a stub runner
    class _Empty:
        def __init__(self, *shape):
            self.shape = shape

        def __getitem__(self, idx):
            return self

    real_torch = ref_base.torch
    try:
        for size in ("1.7B", "0.6B"):
            mc = qwen3_tts_config(size)
            for stack in (mc.talker, mc.predictor):
                hf = types.SimpleNamespace(num_key_value_heads=stack.num_key_value_heads, head_dim=stack.head_dim,
                                           hidden_size=stack.hidden_size, num_attention_heads=stack.num_attention_heads,
                                           num_hidden_layers=stack.num_hidden_layers)
                bb = mem.kv_block_bytes(stack)
                for total, free, alloc, peak, cur, util, frac in (
                        (191_514_542_080, 180 << 30, 4 << 30, 6 << 30, 4 << 30, 0.9, None),
                        (191_514_542_080, 180 << 30, 4 << 30, 6 << 30, 4 << 30, 0.4278, 0.9),
                        (191_514_542_080, 100 << 30, 9 << 30, 12 << 30, 9 << 30, 0.85, 0.5),
                        (85_899_345_920, 70 << 30, 5 << 30, 5 << 30, 5 << 30, 0.3, None),
                        (25_769_803_776, 2 << 30, 3 << 30, 4 << 30, 3 << 30, 0.9, None),       # nothing left: 1 page
                        (25_769_803_776, 20 << 30, 3 << 30, 4 << 30, 3 << 30, 0.05, 0.3)):
                    cfg = types.SimpleNamespace(gpu_memory_utilization=util, process_gpu_memory_fraction=frac,
                                                num_kvcache_blocks=-1)
                    ref_base.torch = types.SimpleNamespace(
                        bfloat16=real_torch.bfloat16, empty=_Empty,
                        cuda=types.SimpleNamespace(
                            mem_get_info=lambda: (free, total), memory_allocated=lambda: alloc,
                            memory_stats=lambda: {"allocated_bytes.all.peak": peak, "allocated_bytes.all.current": cur}))
                    runner = types.SimpleNamespace(config=cfg, model_config=hf, world_size=1, block_size=256,
                                                   model=types.SimpleNamespace(modules=lambda: []))
                    import warnings
                    with warnings.catch_warnings():
                        warnings.simplefilter("ignore")
                        ref_base.ModelRunner.allocate_kv_cache(runner)
                    got, _ = mem.num_kv_blocks(total, free, alloc, peak, cur, util, frac, bb)
                    assert got == cfg.num_kvcache_blocks, (size, total, free, util, frac, got, cfg.num_kvcache_blocks)
                    rows["blocks"].append(dict(model=size, layers=stack.num_hidden_layers, total=total, free=free,
                                               allocated=alloc, peak=peak, current=cur, gpu_memory_utilization=util,
                                               process_gpu_memory_fraction=frac, block_bytes=bb,
                                               num_kvcache_blocks=cfg.num_kvcache_blocks))
    finally:
        ref_base.torch = real_torch
    return rows


def prompt_config():
    """Token ids of the prompt assembly (config.py of the reference: Qwen3TTSConfig / talker_config), toy values."""
    tc = types.SimpleNamespace(
        num_code_groups=16, codec_nothink_id=5, codec_think_id=6, codec_think_bos_id=7, codec_think_eos_id=8,
        codec_pad_id=9, codec_bos_id=10, codec_language_id={"english": 20, "chinese": 21, "sichuan": 22},
        spk_id={"anna": 30, "li": 31}, spk_is_dialect={"li": "sichuan"})
    return types.SimpleNamespace(talker_config=tc, tts_bos_token_id=3, tts_eos_token_id=4, tts_pad_token_id=2)


def prompt_cases(g):
    """Seeded requests covering every branch of utils/generation.py:18-379, 412-515."""
    ids = lambda n: torch.randint(11, 290, (1, n), generator=g)
    code = lambda t: torch.randint(0, 90, (t, 16), generator=g)
    vec = lambda: randn(g, 48, std=0.1)
    clone = lambda codes, xv, icl: {"ref_code": codes, "ref_spk_embedding": [vec() for _ in codes],
                                    "x_vector_only_mode": xv, "icl_mode": icl}
    return {
        "plain_stream": dict(input_ids=[ids(14)]),
        "plain_nonstream": dict(input_ids=[ids(14)], non_streaming_mode=True),
        "instruct_speaker_language": dict(input_ids=[ids(12)], instruct_ids=[ids(6)[0]], speakers=["Anna"], languages=["English"]),
        "dialect": dict(input_ids=[ids(11)], speakers=["li"], languages=["auto"]),
        "batch_mixed": dict(input_ids=[ids(10), ids(17), ids(13)], instruct_ids=[None, ids(4), None],
                            speakers=[None, "anna", ""], languages=["auto", "chinese", "english"]),
        "batch_nonstream": dict(input_ids=[ids(16), ids(11)], speakers=["li", None], languages=["chinese", "auto"],
                                non_streaming_mode=True),
        "xvector": dict(input_ids=[ids(12)], voice_clone_prompt=clone([None], [True], [False]), languages=["english"]),
        "icl_text_longer": dict(input_ids=[ids(40)], ref_ids=[ids(12)], voice_clone_prompt=clone([code(9)], [False], [True])),
        "icl_codec_longer": dict(input_ids=[ids(12)], ref_ids=[ids(9)], voice_clone_prompt=clone([code(30)], [False], [True]),
                                 languages=["english"]),
        "icl_nonstream": dict(input_ids=[ids(15)], ref_ids=[ids(10)], voice_clone_prompt=clone([code(12)], [False], [True]),
                              non_streaming_mode=True),
    }


def pin_prompt():
    """oracle.prompt == the live utils/generation.py (bit-exact, CPU bf16), every mode."""
    from oracle import prompt as op
    from nano_qwen3tts_vllm.utils import generation as ref
    from nano_qwen3tts_vllm.models.qwen3_tts_talker import Qwen3TTSTalkerResizeMLP
    g = torch.Generator().manual_seed(20261105)
    cfg = prompt_config()
    text_emb = torch.nn.Embedding(300, 64).to(BF16)
    codec_emb = torch.nn.Embedding(200, 48).to(BF16)
    proj = Qwen3TTSTalkerResizeMLP(64, 64, 48, "silu", bias=True).to(BF16)
    books = [torch.nn.Embedding(100, 48).to(BF16) for _ in range(15)]
    with torch.no_grad():
        for m in [text_emb, codec_emb, *books]:
            m.weight.copy_(randn(g, *m.weight.shape, std=0.05))
        for lin in (proj.linear_fc1, proj.linear_fc2):
            lin.weight.copy_(randn(g, *lin.weight.shape, std=0.12))
            lin.bias.copy_(randn(g, *lin.bias.shape, std=0.02))
    out = {"weights": {"text_embedding": text_emb.weight.data.clone(), "codec_embedding": codec_emb.weight.data.clone(),
                       "fc1_w": proj.linear_fc1.weight.data.clone(), "fc1_b": proj.linear_fc1.bias.data.clone(),
                       "fc2_w": proj.linear_fc2.weight.data.clone(), "fc2_b": proj.linear_fc2.bias.data.clone(),
                       "books": torch.stack([b.weight.data for b in books])},
           "cases": {}}

    def icl_fn(text_id, ref_id, ref_code, tts_pad_embed, tts_eos_embed, non_streaming_mode):
        return ref.generate_icl_prompt(text_id=text_id, ref_id=ref_id, ref_code=ref_code, tts_pad_embed=tts_pad_embed,
                                       tts_eos_embed=tts_eos_embed, non_streaming_mode=non_streaming_mode, config=cfg,
                                       text_embedding=text_emb, input_embedding=codec_emb, text_projection=proj,
                                       code_predictor_embeddings=books, device="cpu")

    with torch.no_grad():
        for name, kw in prompt_cases(g).items():
            want = ref.prepare_inputs(cfg, kw["input_ids"], instruct_ids=kw.get("instruct_ids"), ref_ids=kw.get("ref_ids"),
                                      voice_clone_prompt=kw.get("voice_clone_prompt"), languages=kw.get("languages"),
                                      speakers=kw.get("speakers"), non_streaming_mode=kw.get("non_streaming_mode", False),
                                      text_embedding=text_emb, input_embedding=codec_emb, text_projection=proj, device="cpu",
                                      generate_speaker_prompt_fn=lambda v: ref.generate_speaker_prompt(v, "cpu"),
                                      generate_icl_prompt_fn=icl_fn)
            got = op.prepare_inputs(cfg, kw["input_ids"], instruct_ids=kw.get("instruct_ids"), ref_ids=kw.get("ref_ids"),
                                    voice_clone_prompt=kw.get("voice_clone_prompt"), languages=kw.get("languages"),
                                    speakers=kw.get("speakers"), non_streaming_mode=kw.get("non_streaming_mode", False),
                                    text_embedding=text_emb, input_embedding=codec_emb, text_projection=proj,
                                    predictor_embeddings=books)
            for a, b, what in zip(got, want, ("prompts", "trailing", "tts_pad", "mask")):
                eq(a, b, f"prompt {name}: {what}")
            out["cases"][name] = {"inputs": kw, "prompts": want[0], "trailing": want[1], "tts_pad": want[2], "mask": want[3]}
    return out


def main():
    import_reference()
    os.makedirs(GOLDEN, exist_ok=True)
    import json
    policy = pin_memory_policy()
    with open(os.path.join(GOLDEN, "memory_policy.json"), "w") as f:
        json.dump(policy, f, indent=1)
    print(f"memory policy pinned: {len(policy['split'])} split rows, {len(policy['blocks'])} allocation rows")
    ops_vec = pin_ops()
    torch.save(ops_vec, os.path.join(GOLDEN, "ops.pt"))
    print("ops pinned:", ", ".join(ops_vec))
    model_vec = pin_models()
    torch.save(model_vec, os.path.join(GOLDEN, "models.pt"))
    print("models pinned:", ", ".join(k for k in model_vec))
    prompt_vec = pin_prompt()
    torch.save(prompt_vec, os.path.join(GOLDEN, "prompt.pt"))
    print("prompt assembly pinned:", ", ".join(prompt_vec["cases"]))
    for f in sorted(os.listdir(GOLDEN)):
        print(f, os.path.getsize(os.path.join(GOLDEN, f)), "bytes")


if __name__ == "__main__":
    main()
