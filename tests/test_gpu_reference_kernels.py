"""a1 / a2 / a3 / a5 against what the reference EXECUTES on a GPU (these ops are CUDA-only
in the reference, so the CPU oracle's versions are restatements - this file pins them on the box):

  a1  store_kvcache         the reference launches a Triton kernel, one program per token, that copies the token's D
                            contiguous K and V elements to cache[slot*D : slot*D+D] and returns at slot == -1
                            (layers/attention.py:10-30).  `_ref_store_kvcache` below is a Triton kernel written to
                            that description (triton 3.6.0, the reference's own JIT); cache BYTES must be equal.
  a2  decode attention      flash_attn_with_kvcache(q.unsqueeze(1), k_cache, v_cache, cache_seqlens=, block_table=,
                            softmax_scale=, causal=True)                    layers/attention.py:92-95
  a3  prefill attention     flash_attn_varlen_func(q, k, v, cu_seqlens_q/k, max_seqlen_q/k, softmax_scale, causal=True)
                            at the BASELINE shapes: [2] x 64 (predictor), [40], [64] x 64, ragged, [4096] (configs[3])
                                                                               layers/attention.py:78-91
  a5  RoPE                  the reference's eager torch ops ON THE GPU (CUDA cos/sin, bf16 arithmetic):
                            layers/rotary_embedding.py:101-148, 172-212 via oracle.ops run on CUDA tensors - the plain
                            RoPE surface must be bit-identical, and so must the K rows the fused kernel writes to the
                            cache from already-normalised heads.

Tolerances: bytes for a1 / a5; 2e-2 relative (north_star) for attention outputs, flash-attn being the reference."""
import pytest
import torch

from oracle import ops as O

pytestmark = pytest.mark.gpu
BF16 = torch.bfloat16
H, KVH, HD, PAGE = 16, 8, 128, 256
TOL = 2e-2


def rel(a, b):
    a, b = a.float(), b.float()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-12))


# ---------------------------------------------------------------------------------------------------- a1
def _ref_store_kvcache(key, value, k_cache, v_cache, slot_mapping):
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

    n, heads, hd = key.shape
    assert key.stride(-1) == 1 and key.stride(1) == hd and k_cache.stride(1) == heads * hd      # attention.py:49-53
    kernel[(n,)](key, key.stride(0), value, value.stride(0), k_cache, v_cache, slot_mapping, heads * hd)


@pytest.mark.parametrize("n,strided", [(1, False), (64, False), (64, True), (300, True), (4096, False)])
def test_store_kvcache_bytes_equal_triton_reference(native, n, strided):
    pytest.importorskip("triton")
    g = torch.Generator(device="cuda").manual_seed(n)
    npages = (n + PAGE - 1) // PAGE + 2
    if strided:                          # K and V as slices of one QKV projection (row stride != D)
        qkv = torch.randn(n, (H + 2 * KVH) * HD, device="cuda", generator=g).to(BF16)
        k = qkv[:, H * HD:(H + KVH) * HD].view(n, KVH, HD)
        v = qkv[:, (H + KVH) * HD:].view(n, KVH, HD)
    else:
        k = torch.randn(n, KVH, HD, device="cuda", generator=g).to(BF16)
        v = torch.randn(n, KVH, HD, device="cuda", generator=g).to(BF16)
    slots = torch.randperm(npages * PAGE, device="cuda", generator=g)[:n].to(torch.int32)
    if n > 2:
        slots[1] = -1                    # graph padding rows are skipped
        slots[n // 2] = -1
    base_k = torch.randn(npages, PAGE, KVH, HD, device="cuda", generator=g).to(BF16)
    base_v = torch.randn(npages, PAGE, KVH, HD, device="cuda", generator=g).to(BF16)
    kc_ref, vc_ref, kc, vc = base_k.clone(), base_v.clone(), base_k.clone(), base_v.clone()
    _ref_store_kvcache(k, v, kc_ref, vc_ref, slots)
    native.store_kvcache(k, v, kc, vc, slots)
    torch.cuda.synchronize()
    assert torch.equal(kc.view(torch.int16), kc_ref.view(torch.int16))
    assert torch.equal(vc.view(torch.int16), vc_ref.view(torch.int16))
    # the fused q/k-norm + RoPE + write kernel: V rows are raw copies, K rows are its own k_out
    kc2, vc2 = base_k.clone(), base_v.clone()
    pos = torch.arange(n, device="cuda", dtype=torch.int64)
    inv_freq = O.make_inv_freq(HD, 1e6).cuda()
    q = torch.randn(n, H, HD, device="cuda", generator=g).to(BF16)
    w = torch.ones(HD, device="cuda", dtype=BF16)
    _, k_out = native.qknorm_rope_kvwrite(q, k, v, w, w, 1e-6, pos, inv_freq, slot_mapping=slots, k_cache=kc2, v_cache=vc2)
    kc3, vc3 = base_k.clone(), base_v.clone()
    _ref_store_kvcache(k_out, v, kc3, vc3, slots)
    torch.cuda.synchronize()
    assert torch.equal(kc2.view(torch.int16), kc3.view(torch.int16))
    assert torch.equal(vc2.view(torch.int16), vc3.view(torch.int16))


# ---------------------------------------------------------------------------------------------------- a3
@pytest.mark.parametrize("lens", [[2] * 64, [40], [64] * 64, [300, 5], [1, 17, 256, 257], [4096]], ids=str)
def test_varlen_prefill_vs_flash_attn_varlen_func(native, lens):
    fa = pytest.importorskip("flash_attn")
    g = torch.Generator(device="cuda").manual_seed(sum(lens))
    t = sum(lens)
    qkv = torch.randn(t, (H + 2 * KVH) * HD, device="cuda", generator=g).to(BF16)      # slices of one projection
    q = qkv[:, :H * HD].view(t, H, HD)
    k = qkv[:, H * HD:(H + KVH) * HD].view(t, KVH, HD)
    v = qkv[:, (H + KVH) * HD:].view(t, KVH, HD)
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32, device="cuda")
    scale = HD ** -0.5
    want = fa.flash_attn_varlen_func(q, k, v, cu_seqlens_q=cu, cu_seqlens_k=cu, max_seqlen_q=max(lens),
                                     max_seqlen_k=max(lens), softmax_scale=scale, causal=True)     # attention.py:88-91
    got = native.varlen_attn_prefill(q, k, v, cu, cu, max(lens), max(lens), scale)
    torch.cuda.synchronize()
    assert rel(got, want) <= TOL, f"lens={lens[:4]}...: rel err {rel(got, want):.4g}"
    if t <= 512:                                                                      # and the CPU oracle agrees with both
        ora = O.varlen_attn_prefill(q.cpu(), k.cpu(), v.cpu(), cu.cpu(), cu.cpu(), scale)
        assert rel(got.cpu(), ora) <= TOL and rel(want.cpu(), ora) <= TOL


# ---------------------------------------------------------------------------------------------------- a2
@pytest.mark.parametrize("b,max_len", [(1, 577), (8, 300), (64, 600), (8, 4352), (64, 17)])
def test_paged_decode_vs_flash_attn_with_kvcache(native, b, max_len):
    fa = pytest.importorskip("flash_attn")
    g = torch.Generator(device="cuda").manual_seed(b * 7 + max_len)
    lens = torch.randint(1, max_len + 1, (b,), generator=g, device="cuda", dtype=torch.int32)
    lens[0] = max_len
    width = (max_len + PAGE - 1) // PAGE
    npages = b * width + 1
    bt = torch.randperm(npages, generator=g, device="cuda")[: b * width].to(torch.int32).view(b, width).contiguous()
    kc = torch.randn(npages, PAGE, KVH, HD, device="cuda", generator=g).to(BF16)
    vc = torch.randn(npages, PAGE, KVH, HD, device="cuda", generator=g).to(BF16)
    q = torch.randn(b, H, HD, device="cuda", generator=g).to(BF16)
    scale = HD ** -0.5
    want = fa.flash_attn_with_kvcache(q.unsqueeze(1), kc, vc, cache_seqlens=lens, block_table=bt, softmax_scale=scale,
                                      causal=True).squeeze(1)                          # attention.py:93-95
    got = native.paged_attn_decode(q, kc, vc, bt, lens, scale, width * PAGE)
    torch.cuda.synchronize()
    assert rel(got, want) <= TOL


# ---------------------------------------------------------------------------------------------------- a5
@pytest.mark.parametrize("mrope", [None, "interleaved", "sectioned"])
def test_rope_bit_exact_vs_reference_ops_on_cuda(native, mrope):
    """The reference's RoPE is eager torch; on the GPU box its cos/sin are CUDA's.  oracle.ops.rope is that op
    sequence (pinned bit-exactly against rotary_embedding.py on CPU): run on CUDA tensors it IS the reference's GPU
    arithmetic, and the kernel must reproduce it to the bit - q, k and the K rows written to the cache."""
    g = torch.Generator(device="cuda").manual_seed(11)
    n = 257
    q = torch.randn(n, H, HD, device="cuda", generator=g).to(BF16)
    k = torch.randn(n, KVH, HD, device="cuda", generator=g).to(BF16)
    v = torch.randn(n, KVH, HD, device="cuda", generator=g).to(BF16)
    pos = torch.randint(0, 4400, (n,), device="cuda", generator=g, dtype=torch.int64)
    pos[:4] = torch.tensor([0, 1, 4095, 4351], device="cuda")
    inv_freq = O.make_inv_freq(HD, 1e6).cuda()
    modality, mod_t, positions = None, None, pos
    if mrope:
        modality = O.mrope_dim_modality(HD, [24, 20, 20], mrope == "interleaved")
        mod_t = torch.tensor(modality, dtype=torch.int32, device="cuda")
        positions = torch.stack([pos, pos // 2 + 3, pos // 3 + 11]).contiguous()
    want_q, want_k = O.rope(positions, q, k, inv_freq, 1.0, modality)                 # torch CUDA ops
    slots = torch.randperm(2 * PAGE, device="cuda", generator=g)[:n].to(torch.int32)
    kc = torch.zeros(2, PAGE, KVH, HD, device="cuda", dtype=BF16)
    vc = torch.zeros_like(kc)
    got_q, got_k = native.qknorm_rope_kvwrite(q, k, v, None, None, 1e-6, positions, inv_freq, modality=mod_t,
                                              slot_mapping=slots, k_cache=kc, v_cache=vc)
    torch.cuda.synchronize()
    assert torch.equal(got_q.view(torch.int16), want_q.view(torch.int16)), \
        f"{(got_q != want_q).sum().item()} of {got_q.numel()} q elements differ"
    assert torch.equal(got_k.view(torch.int16), want_k.view(torch.int16))
    kflat, vflat = kc.view(-1, KVH * HD), vc.view(-1, KVH * HD)
    assert torch.equal(kflat[slots.long()].view(torch.int16), want_k.reshape(n, -1).view(torch.int16))     # KV slots: bit-exact
    assert torch.equal(vflat[slots.long()].view(torch.int16), v.reshape(n, -1).view(torch.int16))


@pytest.mark.parametrize("rounding", [0, 1])
def test_fused_qknorm_rope_kv_slots_vs_reference_ops_on_cuda(native, rounding):
    """q_norm / k_norm + RoPE + KV write (models/qwen3_tts_share.py:96-106) against the same ops in torch on CUDA.
    The rotation is bit-exact (test above); the head norm's mean over 128 values is summed in a different order than
    torch's CUDA reduction, so rstd can differ in its last fp32 bit and the bf16 rounding of a NORMALISED value can
    flip by one ulp; the rotation then mixes that value into two outputs (x*cos + rotate_half(x)*sin), so an affected
    output moves by at most ~2 ulp of the head's largest element.  A tiny share of the elements, everything else
    byte-identical."""
    g = torch.Generator(device="cuda").manual_seed(12 + rounding)
    n = 200
    q = torch.randn(n, H, HD, device="cuda", generator=g).to(BF16)
    k = torch.randn(n, KVH, HD, device="cuda", generator=g).to(BF16)
    v = torch.randn(n, KVH, HD, device="cuda", generator=g).to(BF16)
    qw = (1 + 0.1 * torch.randn(HD, device="cuda", generator=g)).to(BF16)
    kw = (1 + 0.1 * torch.randn(HD, device="cuda", generator=g)).to(BF16)
    pos = torch.randint(0, 4400, (n,), device="cuda", generator=g, dtype=torch.int64)
    inv_freq = O.make_inv_freq(HD, 1e6).cuda()
    want_q, want_k = O.qknorm_rope(q, k, qw, kw, 1e-6, pos, inv_freq, fused_rounding=bool(rounding))
    slots = torch.arange(n, device="cuda", dtype=torch.int32)
    kc = torch.zeros(1, PAGE, KVH, HD, device="cuda", dtype=BF16)
    vc = torch.zeros_like(kc)
    got_q, _ = native.qknorm_rope_kvwrite(q, k, v, qw, kw, 1e-6, pos, inv_freq, slot_mapping=slots, k_cache=kc, v_cache=vc,
                                          want_k_out=False, rounding=rounding)
    torch.cuda.synchronize()
    got_k = kc.view(-1, KVH, HD)[:n]
    for name, got, want in (("q", got_q, want_q), ("k_cache", got_k, want_k)):
        ne = got != want
        frac = ne.float().mean().item()
        assert frac <= 5e-3, f"{name}: {frac:.3%} of the elements differ"
        if ne.any():
            head_max = want.float().abs().amax(dim=-1, keepdim=True).expand_as(want)
            ulp = torch.exp2(torch.floor(torch.log2(head_max.clamp_min(1e-30))) - 7)
            worst = ((got.float() - want.float()).abs()[ne] / ulp[ne]).max().item()
            assert worst <= 2.01, f"{name}: {worst:.2f} ulp of the head maximum"
    assert torch.equal(vc.view(-1, KVH, HD)[:n].view(torch.int16), v.view(torch.int16))
