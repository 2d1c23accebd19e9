"""GPU parity of the persistent dataflow decoder stack (b2tts_decoder_stack_decode, csrc/decoder_stack.cu) against
the separate launches it replaces, layer by layer (models/qwen3_tts_share.py:163-176 of the reference):
b2tts_rmsnorm -> b2tts_gemm_bf16 -> b2tts_fused_decode_attn -> b2tts_gemm_bf16(residual) -> b2tts_rmsnorm ->
b2tts_gemm_bf16(SiLU*mul) -> b2tts_gemm_bf16(residual), final b2tts_rmsnorm.  With the same K splits every
rounding point and summation order is the same, so hidden states and KV-cache bytes must be IDENTICAL."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu
BF16 = torch.bfloat16
PAGE = 256


def _stack(layers, rounding, hidden=1024, inter=3072, seed=1234):
    from nano_qwen3tts_vllm_b200.config import StackConfig
    from nano_qwen3tts_vllm_b200.models.qwen3_tts_share import DecoderStack, random_init_
    cfg = StackConfig(hidden_size=hidden, intermediate_size=inter, num_hidden_layers=layers, num_attention_heads=16,
                      num_key_value_heads=8, head_dim=128, vocab_size=2048)
    torch.set_default_dtype(BF16)
    try:
        stack = DecoderStack(cfg, rounding).to("cuda")
    finally:
        torch.set_default_dtype(torch.float32)
    random_init_(stack, seed)
    return cfg, stack


def _layout(b, max_len, seed, pad_rows=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    dev = "cuda"
    lens = torch.randint(1, max_len + 1, (b,), generator=g, device=dev, dtype=torch.int32)
    lens[0] = max_len
    if b > 1:
        lens[1] = 1
    width = (max_len + PAGE - 1) // PAGE
    npages = b * width + 2
    perm = torch.randperm(npages, generator=g, device=dev)[: b * width].to(torch.int32)
    bt = perm.view(b, width).contiguous()
    slots = (bt.gather(1, ((lens - 1) // PAGE).long().view(b, 1)).view(-1) * PAGE + (lens - 1) % PAGE).to(torch.int32)
    if pad_rows:
        lens[-pad_rows:] = 0
        slots[-pad_rows:] = -1
    pos = (lens.long() - 1).clamp_min(0).contiguous()
    return lens, bt, slots, pos, npages, width


def _reference(nat, stack, cfg, x, pos, slots, bt, lens, caches, plan, max_ctx):
    """The separate launches with the dataflow kernel's K splits."""
    from nano_qwen3tts_vllm_b200.layers.linear import packed_weight
    so, sd = plan
    eps, rnd = cfg.rms_norm_eps, stack.rounding
    b = x.shape[0]
    resid = x.clone()
    for layer, (kc, vc) in zip(stack.layers, caches):
        a, mlp = layer.self_attn, layer.mlp
        h = nat.rmsnorm(resid, layer.input_layernorm.weight, eps, rnd)
        qkv = nat.gemm_bf16(h, packed_weight(a.qkv_proj.weight), a.qkv_proj.weight.shape[0], splits=1)
        q = qkv[:, :cfg.q_size].view(b, 16, 128)
        k = qkv[:, cfg.q_size:cfg.q_size + cfg.kv_size].view(b, 8, 128)
        v = qkv[:, cfg.q_size + cfg.kv_size:].view(b, 8, 128)
        ws = nat.fused_decode_attn_workspace(b, 16, 128, max_ctx, "cuda")
        o = nat.fused_decode_attn(q, k, v, a.q_norm.weight, a.k_norm.weight, eps, pos, a.rotary_emb.inv_freq, slots, kc, vc,
                                  bt, lens, a.scaling, max_ctx, ws, rounding=rnd).view(b, cfg.q_size)
        resid = nat.gemm_bf16(o, packed_weight(a.o_proj.weight), cfg.hidden_size, residual=resid,
                              epilogue=nat.GEMM_RESIDUAL, splits=so)
        h = nat.rmsnorm(resid, layer.post_attention_layernorm.weight, eps, rnd)
        act = nat.gemm_bf16(h, packed_weight(mlp.gate_up_proj.weight, nat.PACK_GATE_UP), 2 * cfg.intermediate_size,
                            epilogue=nat.GEMM_SILU_MUL, rounding=rnd, splits=1)
        resid = nat.gemm_bf16(act, packed_weight(mlp.down_proj.weight), cfg.hidden_size, residual=resid,
                              epilogue=nat.GEMM_RESIDUAL, splits=sd)
    return nat.rmsnorm(resid, stack.norm.weight, eps, rnd)


def _run_case(nat, b, max_len, layers, rounding, seed, pad_rows=0, repeat=1):
    from nano_qwen3tts_vllm_b200.models import qwen3_tts_share as share
    from nano_qwen3tts_vllm_b200.utils.context import reset_context, set_context
    share.DATAFLOW_STACK = True                     # opt-in path (off by default)
    cfg, stack = _stack(layers, rounding)
    plan = stack.dataflow_plan()
    assert plan is not None, "predictor-sized stack must fit the dataflow kernel on this GPU"
    lens, bt, slots, pos, npages, width = _layout(b, max_len, seed, pad_rows)
    g = torch.Generator(device="cuda").manual_seed(seed + 1)
    base = [(torch.randn(npages, PAGE, 8, 128, device="cuda", generator=g).to(BF16),
             torch.randn(npages, PAGE, 8, 128, device="cuda", generator=g).to(BF16)) for _ in range(layers)]
    x = (torch.randn(b, cfg.hidden_size, device="cuda", generator=g) * 0.5).to(BF16)
    max_ctx = width * PAGE
    os.environ["B2TTS_FA_NSPLIT"] = "1"            # the reference attention without split-KV: same summation order
    try:
        ref_caches = [(k.clone(), v.clone()) for k, v in base]
        want = _reference(nat, stack, cfg, x, pos, slots, bt, lens, ref_caches, plan, max_ctx)
    finally:
        del os.environ["B2TTS_FA_NSPLIT"]
    for rep in range(repeat):
        caches = [(k.clone(), v.clone()) for k, v in base]
        for layer, (kc, vc) in zip(stack.layers, caches):
            layer.self_attn.attn.k_cache, layer.self_attn.attn.v_cache = kc, vc
        set_context(False, slot_mapping=slots, context_lens=lens, block_tables=bt, max_context_len=max_ctx)
        try:
            x_before = x.clone()
            got = stack.run_layers(x, pos)
            torch.cuda.synchronize()
        finally:
            reset_context()
        assert torch.equal(x, x_before), "the input rows were modified"
        live = b - pad_rows
        for li, ((k1, v1), (k2, v2)) in enumerate(zip(ref_caches, caches)):
            assert torch.equal(k1, k2), f"layer {li}: K cache bytes differ (rep {rep})"
            assert torch.equal(v1, v2), f"layer {li}: V cache bytes differ (rep {rep})"
        diff = (got[:live].float() - want[:live].float()).abs().max().item()
        assert torch.equal(got[:live], want[:live]), f"hidden states differ: max abs {diff} (rep {rep})"


@pytest.mark.parametrize("b,max_len,layers", [(1, 3, 1), (8, 17, 2), (64, 17, 5), (33, 300, 5), (64, 16, 5), (16, 700, 3)])
def test_dataflow_stack_matches_separate_launches(native, b, max_len, layers):
    _run_case(native, b, max_len, layers, 0, seed=b * 100 + layers)


def test_dataflow_stack_fused_rounding_padding_rows_and_replay(native):
    """ROUND_FUSED, CUDA-graph padding rows (context 0, slot -1), and repeated launches on one workspace
    (the arrival counters are epoch-scaled: no reset between launches)."""
    _run_case(native, 16, 40, 5, 1, seed=7, pad_rows=3, repeat=3)
    _run_case(native, 64, 17, 5, 0, seed=8, repeat=4)


def test_dataflow_stack_in_cuda_graph(native):
    """Captured once, replayed many times: identical output every time."""
    from nano_qwen3tts_vllm_b200.models import qwen3_tts_share as share
    from nano_qwen3tts_vllm_b200.utils.context import reset_context, set_context
    share.DATAFLOW_STACK = True
    cfg, stack = _stack(5, 0)
    lens, bt, slots, pos, npages, width = _layout(64, 17, 11)
    g = torch.Generator(device="cuda").manual_seed(12)
    for layer in stack.layers:
        layer.self_attn.attn.k_cache = torch.randn(npages, PAGE, 8, 128, device="cuda", generator=g).to(BF16)
        layer.self_attn.attn.v_cache = torch.randn(npages, PAGE, 8, 128, device="cuda", generator=g).to(BF16)
    x = (torch.randn(64, cfg.hidden_size, device="cuda", generator=g) * 0.5).to(BF16)
    set_context(False, slot_mapping=slots, context_lens=lens, block_tables=bt, max_context_len=width * PAGE)
    try:
        eager = stack.run_layers(x, pos).clone()
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            out = stack.run_layers(x, pos)
        for _ in range(5):
            graph.replay()
            torch.cuda.synchronize()
            assert torch.equal(out, eager)
    finally:
        reset_context()


@pytest.fixture(autouse=True)
def _restore_default_path():
    from nano_qwen3tts_vllm_b200.models import qwen3_tts_share as share
    before = share.DATAFLOW_STACK
    yield
    share.DATAFLOW_STACK = before
