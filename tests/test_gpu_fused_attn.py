"""GPU parity of the fused decode attention (b2tts_fused_decode_attn) against the separate launches it
replaces - b2tts_qknorm_rope_kvwrite then b2tts_paged_attn_decode (models/qwen3_tts_share.py:96-109 of the
reference: q_norm, k_norm, rotary_emb, store_kvcache, flash_attn_with_kvcache).  KV-cache bytes must be
identical; attention outputs within 2e-2 relative (north_star), in practice ~1e-3 (split-KV order only)."""
import pytest
import torch

pytestmark = pytest.mark.gpu
BF16 = torch.bfloat16
H, KVH, HD, PAGE = 16, 8, 128, 256


def _case(native, b, max_len, mrope, rounding, seed, pad_rows=0, ws=None, max_ctx_override=None):
    g = torch.Generator(device="cuda").manual_seed(seed)
    dev = "cuda"
    lens = torch.randint(1, max_len + 1, (b,), generator=g, device=dev, dtype=torch.int32)
    lens[0] = max_len
    if b > 1:
        lens[1] = 1
    width = (max_len + PAGE - 1) // PAGE
    npages = b * width + 3
    perm = torch.randperm(npages, generator=g, device=dev)[: b * width].to(torch.int32)
    bt = perm.view(b, width).contiguous()
    slots = (bt.gather(1, ((lens - 1) // PAGE).long().view(b, 1)).view(-1) * PAGE + (lens - 1) % PAGE).to(torch.int32)
    if pad_rows:                                     # CUDA-graph padding rows: slot -1, context 0
        lens[-pad_rows:] = 0
        slots[-pad_rows:] = -1
    kc = torch.randn(npages, PAGE, KVH, HD, device=dev, generator=g).to(BF16)
    vc = torch.randn(npages, PAGE, KVH, HD, device=dev, generator=g).to(BF16)
    qkv = torch.randn(b, (H + 2 * KVH) * HD, device=dev, generator=g).to(BF16)
    q = qkv[:, :H * HD].view(b, H, HD)
    k = qkv[:, H * HD:(H + KVH) * HD].view(b, KVH, HD)
    v = qkv[:, (H + KVH) * HD:].view(b, KVH, HD)
    qw = (1 + 0.1 * torch.randn(HD, device=dev, generator=g)).to(BF16)
    kw = (1 + 0.1 * torch.randn(HD, device=dev, generator=g)).to(BF16)
    inv_freq = (1.0 / (1e6 ** (torch.arange(0, HD, 2, device=dev).float() / HD))).contiguous()
    if mrope:
        base = (lens.long() - 1).clamp_min(0)
        pos = torch.stack([base, base + 3, base + 7]).contiguous()
        modality = torch.tensor(([0] * 24 + [1] * 20 + [2] * 20) * 2, dtype=torch.int32, device=dev)
    else:
        pos, modality = (lens.long() - 1).clamp_min(0).contiguous(), None
    eps, scale, max_ctx = 1e-6, HD ** -0.5, width * PAGE
    if max_ctx_override:                             # table capacity larger than the longest row: more KV splits
        max_ctx = max_ctx_override
        bt = torch.cat([bt, torch.full((b, max_ctx // PAGE - width), -1, dtype=torch.int32, device=dev)], 1).contiguous()
    # separate launches
    kc1, vc1 = kc.clone(), vc.clone()
    q_rot, _ = native.qknorm_rope_kvwrite(q, k, v, qw, kw, eps, pos, inv_freq, modality=modality, slot_mapping=slots,
                                          k_cache=kc1, v_cache=vc1, want_k_out=False, rounding=rounding)
    want = native.paged_attn_decode(q_rot, kc1, vc1, bt, lens, scale, max_ctx)
    # one launch
    kc2, vc2 = kc.clone(), vc.clone()
    if ws is None:
        ws = native.fused_decode_attn_workspace(b, H, HD, max_ctx, dev)
    got = native.fused_decode_attn(q, k, v, qw, kw, eps, pos, inv_freq, slots, kc2, vc2, bt, lens, scale, max_ctx, ws,
                                   modality=modality, rounding=rounding)
    got2 = native.fused_decode_attn(q, k, v, qw, kw, eps, pos, inv_freq, slots, kc2, vc2, bt, lens, scale, max_ctx, ws,
                                    modality=modality, rounding=rounding)
    torch.cuda.synchronize()
    assert torch.equal(kc1, kc2), "K cache bytes differ"
    assert torch.equal(vc1, vc2), "V cache bytes differ"
    assert torch.equal(got, got2), "not deterministic / tickets not reset"
    err = (got.float() - want.float()).abs().max().item()
    assert err <= 2e-2 * want.float().abs().max().item() + 1e-6, f"max abs err {err}"
    if pad_rows:
        assert got[-pad_rows:].abs().max().item() == 0
    return err


@pytest.mark.parametrize("b,max_len", [(1, 5), (1, 300), (8, 17), (8, 600), (64, 257), (64, 700), (100, 90), (128, 256)])
@pytest.mark.parametrize("mrope", [False, True])
def test_fused_decode_attention_matches_separate_launches(native, b, max_len, mrope):
    _case(native, b, max_len, mrope, 0, seed=b * 1000 + max_len)


def test_fused_decode_attention_fused_rounding_and_padding_rows(native):
    _case(native, 16, 300, True, 1, seed=3, pad_rows=5)
    _case(native, 64, 64, False, 0, seed=4, pad_rows=1)


def test_one_workspace_serves_every_batch_size(native):
    """One zero-filled workspace sized for the largest batch, reused by launches of shrinking batch sizes with a
    4096-token table capacity (the runner's layout, engine/model_runner/base.py): the split-KV tickets of one batch
    size must not overlap the partial accumulators another batch size wrote (round-1 advisor finding)."""
    ws = native.fused_decode_attn_workspace(16, H, HD, 4096, "cuda")
    for i, b in enumerate((16, 8, 2, 1, 8, 16, 3)):
        _case(native, b, 700, i % 2 == 0, 0, seed=50 + i, ws=ws, max_ctx_override=4096)


@pytest.mark.parametrize("b,max_len,rounding,pad", [(1, 3, 0, 0), (8, 17, 0, 2), (64, 17, 0, 0), (64, 16, 1, 1), (100, 64, 0, 0),
                                                    (33, 40, 1, 0)])
def test_short_context_kernel_is_bit_identical(native, b, max_len, rounding, pad):
    """Block tables of <= 64 tokens (the code predictor) take the one-warp-per-(sequence, kv head) kernel; same
    arithmetic and summation order as the streaming kernel without split-KV, so every byte must agree."""
    import os
    g = torch.Generator(device="cuda").manual_seed(b * 31 + max_len)
    dev = "cuda"
    lens = torch.randint(1, max_len + 1, (b,), generator=g, device=dev, dtype=torch.int32)
    lens[0] = max_len
    npages = b + 2
    bt = torch.randperm(npages, generator=g, device=dev)[:b].to(torch.int32).view(b, 1).contiguous()
    slots = (bt.view(-1) * PAGE + lens - 1).to(torch.int32)
    if pad:
        lens[-pad:] = 0
        slots[-pad:] = -1
    kc = torch.randn(npages, PAGE, KVH, HD, device=dev, generator=g).to(BF16)
    vc = torch.randn(npages, PAGE, KVH, HD, device=dev, generator=g).to(BF16)
    qkv = torch.randn(b, (H + 2 * KVH) * HD, device=dev, generator=g).to(BF16)
    q = qkv[:, :H * HD].view(b, H, HD)
    k = qkv[:, H * HD:(H + KVH) * HD].view(b, KVH, HD)
    v = qkv[:, (H + KVH) * HD:].view(b, KVH, HD)
    qw = (1 + 0.1 * torch.randn(HD, device=dev, generator=g)).to(BF16)
    kw = (1 + 0.1 * torch.randn(HD, device=dev, generator=g)).to(BF16)
    inv_freq = (1.0 / (1e6 ** (torch.arange(0, HD, 2, device=dev).float() / HD))).contiguous()
    pos = (lens.long() - 1).clamp_min(0).contiguous()
    eps, scale = 1e-6, HD ** -0.5
    ws = native.fused_decode_attn_workspace(b, H, HD, PAGE, dev)
    kc1, vc1, kc2, vc2 = kc.clone(), vc.clone(), kc.clone(), vc.clone()
    os.environ["B2TTS_FA_NSPLIT"] = "1"
    try:
        want = native.fused_decode_attn(q, k, v, qw, kw, eps, pos, inv_freq, slots, kc1, vc1, bt, lens, scale, PAGE, ws,
                                        rounding=rounding)
    finally:
        del os.environ["B2TTS_FA_NSPLIT"]
    got = native.fused_decode_attn(q, k, v, qw, kw, eps, pos, inv_freq, slots, kc2, vc2, bt, lens, scale, max_len, ws,
                                   rounding=rounding)
    torch.cuda.synchronize()
    assert torch.equal(kc1, kc2) and torch.equal(vc1, vc2), "KV cache bytes differ"
    assert torch.equal(got, want), f"max abs {(got.float() - want.float()).abs().max().item()}"
