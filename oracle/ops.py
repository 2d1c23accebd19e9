"""Op-level CPU oracle (torch CPU tensors, every rounding point explicit).

TEST INFRASTRUCTURE ONLY - see ``oracle/__init__.py``.

Citations are ``file:line`` under ``/root/reference/nano-qwen3tts-vllm/``.

Pinned against the live reference (``oracle/pin_against_reference.py`` ->
``tests/golden/``): rmsnorm, add_rmsnorm, rope (1-D and M-RoPE), silu, silu*mul,
linear, sampler, lm heads, decoder layer, talker / predictor stacks.
PARITY UNPINNED by a reference run (CUDA-only in the reference): store_kvcache
(Triton kernel) and the two flash-attn calls; restated from
``layers/attention.py:10-30,73-96`` and cross-checked against the flash-attn
wheel and the reference's Triton kernel on the GPU box
(tests/test_gpu_reference_kernels.py).
"""
from __future__ import annotations

import math

import torch

BF16 = torch.bfloat16


# --------------------------------------------------------------------------
# a1  paged KV write                                   layers/attention.py:10-54
# --------------------------------------------------------------------------
def store_kvcache(key, value, k_cache, v_cache, slot_mapping):
    """Scatter rows of ``key``/``value`` ``[N, kv_heads, hd]`` into the paged caches
    ``[blocks, page, kv_heads, hd]`` at flat slot ``slot_mapping[i]``; ``-1`` skips.

    attention.py:21-30: one program per token, ``cache[slot*D : slot*D+D] = row``.
    Bit copy; later tokens overwrite earlier ones that map to the same slot only
    in program order, which the reference never relies on (slots are unique).
    """
    n, kvh, hd = key.shape
    d = kvh * hd
    assert k_cache.stride(1) == d and v_cache.stride(1) == d      # attention.py:52
    assert slot_mapping.numel() == n                              # attention.py:53
    kflat = k_cache.view(-1, d)
    vflat = v_cache.view(-1, d)
    for i, slot in enumerate(slot_mapping.tolist()):
        if slot == -1:                                            # attention.py:23
            continue
        kflat[slot] = key[i].reshape(d)
        vflat[slot] = value[i].reshape(d)


# --------------------------------------------------------------------------
# a2  paged GQA decode attention                       layers/attention.py:92-95
# --------------------------------------------------------------------------
def _sdpa_one(q, k, v, scale, q_offset):
    """q [Lq,H,hd] f32, k/v [Lk,KVH,hd] f32. Causal with bottom-right alignment
    (query i sees keys 0..q_offset+i), GQA h -> h // (H/KVH). fp32 softmax."""
    lq, h, hd = q.shape
    lk, kvh, _ = k.shape
    g = h // kvh
    kk = k.repeat_interleave(g, dim=1)            # [Lk,H,hd]
    vv = v.repeat_interleave(g, dim=1)
    s = torch.einsum("qhd,khd->hqk", q, kk) * scale
    qi = torch.arange(lq).view(1, lq, 1)
    ki = torch.arange(lk).view(1, 1, lk)
    s = s.masked_fill(ki > qi + q_offset, float("-inf"))
    p = torch.softmax(s, dim=-1)
    return torch.einsum("hqk,khd->qhd", p, vv)


def paged_attn_decode(q, k_cache, v_cache, context_lens, block_table, scale):
    """flash_attn_with_kvcache(q.unsqueeze(1), k_cache, v_cache, cache_seqlens,
    block_table, softmax_scale, causal=True) for one query per sequence
    (attention.py:93-95).  Rows with ``context_lens == 0`` (CUDA-graph padding,
    talker_mode_runner.py:97-100) produce zeros.  Output bf16 [B,H,hd]."""
    b, h, hd = q.shape
    page = k_cache.shape[1]
    out = torch.zeros(b, h, hd, dtype=BF16)
    for i in range(b):
        l = int(context_lens[i])
        if l == 0:
            continue
        npages = (l + page - 1) // page
        pages = block_table[i, :npages].long()
        k = k_cache[pages].reshape(-1, *k_cache.shape[2:])[:l].float()
        v = v_cache[pages].reshape(-1, *v_cache.shape[2:])[:l].float()
        o = _sdpa_one(q[i:i + 1].float(), k, v, scale, q_offset=l - 1)
        out[i] = o[0].to(BF16)
    return out


# --------------------------------------------------------------------------
# a3  varlen causal prefill attention                  layers/attention.py:78-91
# --------------------------------------------------------------------------
def varlen_attn_prefill(q, k, v, cu_seqlens_q, cu_seqlens_k, scale,
                        k_cache=None, v_cache=None, block_table=None):
    """flash_attn_varlen_func(causal=True).  Without ``block_table`` K/V are the
    packed projections; with it (prefix-cache hit, attention.py:79-80) K/V come
    from the paged cache and the causal mask is bottom-right aligned."""
    t, h, hd = q.shape
    out = torch.zeros(t, h, hd, dtype=BF16)
    nseq = cu_seqlens_q.numel() - 1
    for s in range(nseq):
        q0, q1 = int(cu_seqlens_q[s]), int(cu_seqlens_q[s + 1])
        k0, k1 = int(cu_seqlens_k[s]), int(cu_seqlens_k[s + 1])
        lq, lk = q1 - q0, k1 - k0
        if lq == 0:
            continue
        if block_table is None:
            ks, vs = k[k0:k1].float(), v[k0:k1].float()
        else:
            page = k_cache.shape[1]
            npages = (lk + page - 1) // page
            pages = block_table[s, :npages].long()
            ks = k_cache[pages].reshape(-1, *k_cache.shape[2:])[:lk].float()
            vs = v_cache[pages].reshape(-1, *v_cache.shape[2:])[:lk].float()
        o = _sdpa_one(q[q0:q1].float(), ks, vs, scale, q_offset=lk - lq)
        out[q0:q1] = o.to(BF16)
    return out


# --------------------------------------------------------------------------
# a4  RMSNorm                                           layers/layernorm.py:18-41
# --------------------------------------------------------------------------
def rmsnorm(x, weight, eps, fused_rounding=False):
    """layernorm.py:22-27 in eager mode: fp32 normalise, round to bf16, then a
    bf16*bf16->bf16 multiply with the weight (a second rounding).
    ``fused_rounding``: the same source compiled by inductor on a GPU
    (``@torch.compile``, layernorm.py:17) - fp32 through to the one store rounding."""
    xf = x.float()
    var = xf.pow(2).mean(dim=-1, keepdim=True)
    y = xf * torch.rsqrt(var + eps)
    if fused_rounding:
        return (y * weight.float()).to(x.dtype)
    return y.to(x.dtype) * weight


def add_rmsnorm(x, residual, weight, eps):
    """layernorm.py:35-41: the sum is kept in fp32 for the norm; only the
    returned residual is rounded."""
    xf = x.float() + residual.float()
    new_residual = xf.to(x.dtype)
    var = xf.pow(2).mean(dim=-1, keepdim=True)
    y = xf * torch.rsqrt(var + eps)
    return y.to(x.dtype) * weight, new_residual


def residual_add_then_rmsnorm(x, residual, weight, eps):
    """What the decoder layer really does (models/qwen3_tts_share.py:171-173):
    a bf16 add, then ``rms_forward`` on the *rounded* sum."""
    new_residual = residual + x
    return rmsnorm(new_residual, weight, eps), new_residual


# --------------------------------------------------------------------------
# a5  RoPE                                      layers/rotary_embedding.py:5-212
# --------------------------------------------------------------------------
def make_inv_freq(rotary_dim, base):
    """rotary_embedding.py:98,169."""
    return 1.0 / (base ** (torch.arange(0, rotary_dim, 2, dtype=torch.float) / rotary_dim))


def mrope_dim_modality(head_dim, mrope_section=None, interleaved=False):
    """Which of the 3 position rows each of the ``head_dim`` cos/sin columns reads.

    rotary_embedding.py:52-72 (interleaved) and :73-80 (sectioned).  Returns a
    list of ``head_dim`` ints in {0,1,2}; all zeros for 1-D RoPE."""
    half = head_dim // 2
    if mrope_section is None:
        return [0] * head_dim
    if interleaved:
        m = [0] * half
        for i, n in enumerate(mrope_section[1:], 1):
            for f in range(i, n * len(mrope_section), len(mrope_section)):
                if f < half:
                    m[f] = i
        return m + m                                   # cat([...]*2), :66-68
    out = []
    for i, n in enumerate(mrope_section):
        out += [i % 3] * (2 * n)                       # sections doubled, :74
    assert len(out) == head_dim
    return out


def rope_cos_sin(positions, inv_freq, dtype=BF16, attention_scaling=1.0, dim_modality=None):
    """cos/sin [T, head_dim] in ``dtype``.  ``positions`` is int64 ``[T]`` or
    ``[3,T]`` (rotary_embedding.py:113-139, 178-200)."""
    hd = 2 * inv_freq.numel()
    if positions.dim() == 1:
        positions = positions.unsqueeze(0).expand(3, -1)
    pos = positions.float()                                      # [3,T]
    inv_freq = inv_freq.to(positions.device)
    freqs = pos[:, :, None] * inv_freq.float()[None, None, :]    # [3,T,half]; K=1 matmul == product
    emb = torch.cat((freqs, freqs), dim=-1)                      # [3,T,hd]
    cos3 = (emb.cos() * attention_scaling).to(dtype)
    sin3 = (emb.sin() * attention_scaling).to(dtype)
    if dim_modality is None:
        dim_modality = [0] * hd
    sel = torch.tensor(dim_modality, dtype=torch.long, device=positions.device)
    idx = sel.view(1, 1, hd).expand(1, cos3.shape[1], hd)
    cos = torch.gather(cos3, 0, idx)[0]
    sin = torch.gather(sin3, 0, idx)[0]
    return cos, sin


def rotate_half(x):
    """rotary_embedding.py:5-9."""
    h = x.shape[-1] // 2
    return torch.cat((-x[..., h:], x[..., :h]), dim=-1)


def apply_rope(x, cos, sin):
    """x [T,heads,hd] bf16; cos/sin [T,hd] bf16.  Three bf16 roundings:
    ``(x*cos) + (rotate_half(x)*sin)`` evaluated with bf16 eager ops
    (rotary_embedding.py:33-34, 82-83)."""
    c = cos.unsqueeze(1)
    s = sin.unsqueeze(1)
    return (x * c) + (rotate_half(x) * s)


def rope(positions, query, key, inv_freq, attention_scaling=1.0, dim_modality=None):
    cos, sin = rope_cos_sin(positions, inv_freq, query.dtype, attention_scaling, dim_modality)
    return apply_rope(query, cos, sin), apply_rope(key, cos, sin)


def qknorm_rope(q, k, q_weight, k_weight, eps, positions, inv_freq,
                attention_scaling=1.0, dim_modality=None, fused_rounding=False):
    """models/qwen3_tts_share.py:96-104: per-head RMSNorm on q and k, then RoPE."""
    qn = rmsnorm(q, q_weight, eps, fused_rounding)
    kn = rmsnorm(k, k_weight, eps, fused_rounding)
    return rope(positions, qn, kn, inv_freq, attention_scaling, dim_modality)


# --------------------------------------------------------------------------
# a6  SiLU and the gated product       layers/activation.py:12-13, share.py:131-135
# --------------------------------------------------------------------------
def silu(x):
    return torch.nn.functional.silu(x)


def silu_and_mul(gate, up, fused_rounding=False):
    """Eager reference: ``silu(gate)`` rounds to bf16, the product rounds again.
    ``fused_rounding`` is what inductor emits on the GPU for the compiled MLP
    (one rounding from fp32)."""
    if fused_rounding:
        g = gate.float()
        return ((g / (1.0 + torch.exp(-g))) * up.float()).to(gate.dtype)
    return torch.nn.functional.silu(gate) * up


# --------------------------------------------------------------------------
# a7  linear                                             layers/linear.py:51,73,150
# --------------------------------------------------------------------------
def linear(x, weight, bias=None):
    return torch.nn.functional.linear(x, weight, bias)


# --------------------------------------------------------------------------
# a8  sampler                                             layers/sampler.py:15-24
# --------------------------------------------------------------------------
def topk_probs(logits, temperatures, top_k=50):
    """sampler.py:16-21: fp32 logits / T, keep ``l >= kth`` (ties kept), softmax."""
    l = logits.float().div_(temperatures.unsqueeze(dim=1))
    kth = torch.topk(l, top_k)[0][..., -1, None]
    l = l.masked_fill(l < kth, -float("inf"))
    return torch.softmax(l, dim=-1)


def sample_from_probs(probs, noise):
    """ATen ``multinomial(probs, 1)`` without replacement is
    ``argmax(probs / q)`` with ``q ~ Exp(1)`` (aten/src/ATen/native/TensorAdvancedIndexing /
    Distributions multinomial fast path).  ``noise`` is that q."""
    return torch.argmax(probs / noise, dim=-1)


def topk_sample(logits, temperatures, noise, top_k=50):
    return sample_from_probs(topk_probs(logits, temperatures, top_k), noise)
