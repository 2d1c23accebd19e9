"""Kernel-level timing legs of bench.py: every launch of ours that the JSON line quotes, and - beside it - the
kernel the reference runs for the same op on the same shapes (``reference_kernels``: cuBLAS through ``F.linear``,
flash-attn 2.8's ``flash_attn_with_kvcache`` / ``flash_attn_varlen_func``, a Triton restatement of the reference's
store kernel).  Library kernels are measured here, never shipped: the product path does not import this file.

All timings: the launches are captured into one CUDA graph (no host gaps), the replay is bracketed by CUDA events
on the launching stream, median of 3 replays after one warm replay.  Every leg works over data far larger than the
126 MB L2 (different weights / caches per launch), so each launch reads cold HBM."""
from __future__ import annotations

import torch

BF16 = torch.bfloat16


def graph_ms(fn, launches: int, iters: int = 3) -> float:
    fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    times = []
    for it in range(iters + 1):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        if it:
            times.append(e0.elapsed_time(e1) / launches)
    del g
    return sorted(times)[len(times) // 2]


def decoder_layers(runner):
    return [m for m in runner.model.modules() if m.__class__.__name__ == "Qwen3TTSDecoderLayer"]


def layer_gemm_bytes(c, m: int) -> float:
    """Algorithmic bytes of the 4 projections of one decoder layer at M rows: 2*(K*N + M*K + M*N) each."""
    w = 2.0 * (c.hidden_size * (c.q_size + 2 * c.kv_size) + c.q_size * c.hidden_size + 3 * c.hidden_size * c.intermediate_size)
    io = 2.0 * m * (c.hidden_size + (c.q_size + 2 * c.kv_size) + c.q_size + c.hidden_size + c.hidden_size
                    + c.intermediate_size + c.intermediate_size + c.hidden_size)
    return w + io


def gemm_stack(runner, m: int, impl: str = "ours"):
    """The 4 projections of every decoder layer of ``runner``'s model at M = m rows over the model's own weights.
    ours: gemm_stream_kernel with the production epilogues (fused RMSNorm in, residual / SiLU*mul out, row sums of
    squares) under programmatic dependent launch.  cublas: the same 4 products as plain ``F.linear`` launches (what the
    reference's layers/linear.py runs; its separate norm / activation / residual kernels are NOT counted).
    Returns (ms per launch, algorithmic bytes per launch, launches)."""
    import torch.nn.functional as F
    import nano_qwen3tts_vllm_b200._native as nat
    from nano_qwen3tts_vllm_b200.layers.linear import RowNorm, linear, linear_residual, linear_silu_mul
    c = runner.stack_config
    dev = runner.device
    layers = decoder_layers(runner)
    resid = (torch.randn(m, c.hidden_size, device=dev) * 0.5).to(BF16)
    attn = (torch.randn(m, c.q_size, device=dev) * 0.5).to(BF16)
    act0 = (torch.randn(m, c.intermediate_size, device=dev) * 0.5).to(BF16)
    eps, rnd = c.rms_norm_eps, nat.ROUND_EAGER

    from nano_qwen3tts_vllm_b200.models.qwen3_tts_share import FUSED_NORM_MAX_ROWS, FUSED_NORM_MIN_ROWS
    # outside this range the product runs the norms as separate launches (not counted here) and plain GEMMs
    wide = not FUSED_NORM_MIN_ROWS <= m <= FUSED_NORM_MAX_ROWS

    def ours_wide():
        x = resid.clone()
        for layer in layers:
            linear(x, layer.self_attn.qkv_proj.weight)
            linear_residual(attn, layer.self_attn.o_proj.weight, x, out=x)
            act = linear_silu_mul(x, layer.mlp.gate_up_proj.weight, rnd)
            linear_residual(act, layer.mlp.down_proj.weight, x, out=x)
        linear_residual(attn, layers[0].self_attn.o_proj.weight, x, out=x)

    def ours():
        if wide:
            return ours_wide()
        x = resid.clone()
        _, ssq = linear_residual(attn, layers[0].self_attn.o_proj.weight, x, out=x, want_ssq=True)
        for layer in layers:
            linear(x, layer.self_attn.qkv_proj.weight, norm=RowNorm(layer.input_layernorm.weight, ssq, eps, rnd))
            _, ssq = linear_residual(attn, layer.self_attn.o_proj.weight, x, out=x, want_ssq=True)
            act = linear_silu_mul(x, layer.mlp.gate_up_proj.weight, rnd,
                                  norm=RowNorm(layer.post_attention_layernorm.weight, ssq, eps, rnd))
            _, ssq = linear_residual(act, layer.mlp.down_proj.weight, x, out=x, want_ssq=True)

    def cublas():
        F.linear(attn, layers[0].self_attn.o_proj.weight)
        for layer in layers:
            F.linear(resid, layer.self_attn.qkv_proj.weight)
            F.linear(attn, layer.self_attn.o_proj.weight)
            F.linear(resid, layer.mlp.gate_up_proj.weight)
            F.linear(act0, layer.mlp.down_proj.weight)
    n = 4 * len(layers) + 1
    ms = graph_ms(ours if impl == "ours" else cublas, n)
    return ms, layer_gemm_bytes(c, m) / 4.0, n


def _paged_caches(b: int, ctx: int, layers: int, device, page=256, kvh=8, hd=128):
    pages_per = (ctx + page - 1) // page
    npages = b * pages_per
    caches = [(torch.randn(npages, page, kvh, hd, device=device, dtype=BF16),
               torch.randn(npages, page, kvh, hd, device=device, dtype=BF16)) for _ in range(layers)]
    bt = torch.randperm(npages, generator=torch.Generator().manual_seed(13)).to(torch.int32).view(b, pages_per).to(device)
    return caches, bt, pages_per * page


def decode_attention(b: int, ctx, device, layers: int = 4, impl: str = "ours", h=16, kvh=8, hd=128, page=256):
    """One decode attention launch per layer over ``layers`` different paged caches; ``ctx`` = an int (uniform) or an
    int32 tensor [b] of context lengths (the new token included).
    ours: fused_decode_attn_kernel = q/k head norm + M-RoPE + KV write of the new token + paged GQA attention + split-KV
    combine in ONE launch.  flash: ``flash_attn_with_kvcache`` alone (layers/attention.py:92-95; the reference's
    separate norm, RoPE and Triton store launches are not counted).  Bytes: sum_b(4096 * L_b + 8192)."""
    import nano_qwen3tts_vllm_b200._native as nat
    lens = torch.full((b,), int(ctx), dtype=torch.int32, device=device) if isinstance(ctx, int) else ctx.to(device, torch.int32)
    lmax = int(lens.max())
    caches, bt, max_ctx = _paged_caches(b, lmax, layers, device, page, kvh, hd)
    slots = (bt.gather(1, ((lens.long() - 1) // page).view(b, 1)).view(b) * page + (lens - 1) % page).to(torch.int32).contiguous()
    qkv = torch.randn(b, (h + 2 * kvh) * hd, device=device, dtype=BF16)
    q, k, v = (qkv[:, :h * hd].view(b, h, hd), qkv[:, h * hd:(h + kvh) * hd].view(b, kvh, hd),
               qkv[:, (h + kvh) * hd:].view(b, kvh, hd))
    alg = float((4096.0 * lens.float() + 8192.0).sum())
    if impl == "ours":
        w = torch.ones(hd, device=device, dtype=BF16)
        inv_freq = (1.0 / (1e6 ** (torch.arange(0, hd, 2, device=device).float() / hd))).contiguous()
        pos = (lens.long() - 1).contiguous()
        ws = nat.fused_decode_attn_workspace(b, h, hd, max_ctx, device)
        out = torch.empty(b, h, hd, device=device, dtype=BF16)

        def fn():
            for kc, vc in caches:
                nat.fused_decode_attn(q, k, v, w, w, 1e-6, pos, inv_freq, slots, kc, vc, bt, lens, hd ** -0.5, max_ctx, ws, out=out)
    else:
        from flash_attn import flash_attn_with_kvcache
        qc = q.contiguous().unsqueeze(1)

        def fn():
            for kc, vc in caches:
                flash_attn_with_kvcache(qc, kc, vc, cache_seqlens=lens, block_table=bt, softmax_scale=hd ** -0.5, causal=True)
    return graph_ms(fn, layers), alg


def prefill_attention(lens: list[int], device, layers: int = 4, impl: str = "ours", h=16, kvh=8, hd=128):
    """Causal varlen prefill attention over packed q/k/v slices of one QKV projection.  ours: prefill_attn_tc_kernel
    (tcgen05, 128 x 128 tiles, TMEM accumulators) for max_seqlen >= 128; flash: ``flash_attn_varlen_func``
    (layers/attention.py:78-91).  Returns (ms per launch, causal flops per launch)."""
    import nano_qwen3tts_vllm_b200._native as nat
    t = sum(lens)
    qkvs = [torch.randn(t, (h + 2 * kvh) * hd, device=device, dtype=BF16) for _ in range(layers)]
    cu = torch.tensor([0] + torch.tensor(lens).cumsum(0).tolist(), dtype=torch.int32, device=device)
    out = torch.empty(t, h, hd, device=device, dtype=BF16)
    scale, mx = hd ** -0.5, max(lens)

    def split(x):
        return (x[:, :h * hd].view(t, h, hd), x[:, h * hd:(h + kvh) * hd].view(t, kvh, hd), x[:, (h + kvh) * hd:].view(t, kvh, hd))
    if impl == "ours":
        def fn():
            for x in qkvs:
                q, k, v = split(x)
                nat.varlen_attn_prefill(q, k, v, cu, cu, mx, mx, scale, out=out)
    else:
        from flash_attn import flash_attn_varlen_func

        def fn():
            for x in qkvs:
                q, k, v = split(x)
                flash_attn_varlen_func(q, k, v, cu_seqlens_q=cu, cu_seqlens_k=cu, max_seqlen_q=mx, max_seqlen_k=mx,
                                       softmax_scale=scale, causal=True)
    flops = float(sum(4.0 * h * hd * (n * (n + 1) / 2.0) for n in lens))      # QK^T + PV over the causal half
    return graph_ms(fn, layers), flops


def kv_store(n: int, device, layers: int = 8, impl: str = "ours", kvh=8, hd=128, page=256):
    """KV-cache write of n new tokens (K and V rows of one QKV projection) to scattered slots.  ours: store_kvcache
    (stand-alone a1 kernel; in the decode frame the write is fused into the attention launch); triton: the reference's
    store kernel restated (layers/attention.py:10-41: one program per token, D = kv_heads * head_dim).
    Bytes: n * D * 2 B * 2 (K, V) * 2 (read + write)."""
    import nano_qwen3tts_vllm_b200._native as nat
    d = kvh * hd
    npages = (n + page - 1) // page + 4
    srcs = [torch.randn(n, 3 * d, device=device, dtype=BF16) for _ in range(layers)]
    caches = [(torch.zeros(npages, page, kvh, hd, device=device, dtype=BF16),
               torch.zeros(npages, page, kvh, hd, device=device, dtype=BF16)) for _ in range(layers)]
    slots = torch.randperm(npages * page, generator=torch.Generator().manual_seed(5))[:n].to(torch.int32).to(device)
    if impl == "ours":
        def fn():
            for x, (kc, vc) in zip(srcs, caches):
                nat.store_kvcache(x[:, d:2 * d].view(n, kvh, hd), x[:, 2 * d:].view(n, kvh, hd), kc, vc, slots)
    else:
        import triton
        import triton.language as tl

        @triton.jit
        def kernel(k_ptr, k_stride, v_ptr, v_stride, kc_ptr, vc_ptr, slots_ptr, D: tl.constexpr):
            tok = tl.program_id(0)
            slot = tl.load(slots_ptr + tok)
            if slot == -1:
                return
            cols = tl.arange(0, D)
            tl.store(kc_ptr + slot * D + cols, tl.load(k_ptr + tok * k_stride + cols))
            tl.store(vc_ptr + slot * D + cols, tl.load(v_ptr + tok * v_stride + cols))

        def fn():
            for x, (kc, vc) in zip(srcs, caches):
                kernel[(n,)](x[:, d:2 * d], 3 * d, x[:, 2 * d:], 3 * d, kc, vc, slots, d)
    return graph_ms(fn, layers), float(n) * d * 2 * 2 * 2


def sampler(b: int, vocab: int, device, top_k: int = 50, reps: int = 16):
    """topk_sample_row_kernel on [b, vocab] bf16 logits (16 different rows sets per graph)."""
    import nano_qwen3tts_vllm_b200._native as nat
    logits = [(torch.randn(b, vocab, device=device) * 2).to(BF16) for _ in range(reps)]
    noise = torch.empty(b, vocab, device=device, dtype=torch.float32).exponential_(1)
    temps = torch.full((b,), 0.9, device=device, dtype=torch.float32)
    out = torch.empty(b, device=device, dtype=torch.int64)

    def fn():
        for lg in logits:
            nat.topk_sample(lg, temps, noise, top_k=top_k, out=out)
    return graph_ms(fn, reps), float(b) * vocab * 2
