"""GPU parity tests: every entry point of libb200tts (through the C ABI) against the CPU oracle
and the committed golden vectors (outputs of the live reference, tests/golden/).

Bars (BASELINE.json north_star): KV slots and sampled indices bit-exact; element-wise ops
bit-exact up to a counted handful of 1-ulp bf16 flips caused by fp32 summation order / libm
ulps (documented per test); attention within 2e-2 relative.
"""
import math

import pytest
import torch

from oracle import ops as O

pytestmark = pytest.mark.gpu
BF16 = torch.bfloat16
DEV = "cuda"


def bf16_ulp_diff(a, b):
    """Distance in bf16 ulps (on the int16 image, sign-magnitude folded)."""
    ai = a.contiguous().view(torch.int16).to(torch.int32)
    bi = b.contiguous().view(torch.int16).to(torch.int32)
    ai = torch.where(ai < 0, -(ai & 0x7FFF), ai)
    bi = torch.where(bi < 0, -(bi & 0x7FFF), bi)
    return (ai - bi).abs()


def assert_bf16_close(got, want, max_ulp=1, max_frac=2e-3, what=""):
    got, want = got.cpu(), want.cpu()
    assert got.shape == want.shape and got.dtype == want.dtype, (what, got.shape, want.shape)
    d = bf16_ulp_diff(got, want)
    nbad = int((d > 0).sum())
    assert int(d.max()) <= max_ulp, f"{what}: max ulp diff {int(d.max())}"
    assert nbad <= max_frac * got.numel() + 1, f"{what}: {nbad}/{got.numel()} elements differ"


def rel_err(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-12))


# ------------------------------------------------------------------------------------------ a1
@pytest.mark.parametrize("n,stride_pad", [(1, 0), (7, 0), (64, 3072), (513, 0)])
def test_store_kvcache_bit_exact(native, n, stride_pad):
    g = torch.Generator().manual_seed(n)
    kvh, hd, page, blocks = 8, 128, 256, 4
    d = kvh * hd
    wide = torch.randn(n, d + stride_pad + d, generator=g).to(BF16)     # k and v slices of one buffer
    key, value = wide[:, :d].view(n, kvh, hd), wide[:, d + stride_pad:].view(n, kvh, hd)
    slots = torch.randperm(blocks * page, generator=g)[:n].to(torch.int32)
    slots[::5] = -1                                                      # skipped rows (attention.py:23)
    kc = torch.randn(blocks, page, kvh, hd, generator=g).to(BF16)
    vc = torch.randn(blocks, page, kvh, hd, generator=g).to(BF16)
    kc_ref, vc_ref = kc.clone(), vc.clone()
    O.store_kvcache(key, value, kc_ref, vc_ref, slots)
    kd, vd = kc.to(DEV), vc.to(DEV)
    wd = wide.to(DEV)
    native.store_kvcache(wd[:, :d].view(n, kvh, hd), wd[:, d + stride_pad:].view(n, kvh, hd), kd, vd, slots.to(DEV))
    assert torch.equal(kd.cpu(), kc_ref) and torch.equal(vd.cpu(), vc_ref)


def test_store_kvcache_empty_and_errors(native):
    kc = torch.zeros(1, 256, 8, 128, dtype=BF16, device=DEV)
    e = torch.zeros(0, 8, 128, dtype=BF16, device=DEV)
    native.store_kvcache(e, e, kc, kc.clone(), torch.zeros(0, dtype=torch.int32, device=DEV))
    with pytest.raises(RuntimeError):
        native.store_kvcache(e.cpu(), e.cpu(), kc, kc, torch.zeros(0, dtype=torch.int32, device=DEV))


# ------------------------------------------------------------------------------------------ a4
@pytest.mark.parametrize("name", ["rms_hidden", "rms_head", "rms_1024"])
def test_rmsnorm_golden(native, golden_ops, name):
    v = golden_ops[name]
    x, res, w = v["x"].to(DEV), v["res"].to(DEV), v["w"].to(DEV)
    y = native.rmsnorm(x, w, v["eps"])
    # fp32 sum-of-squares order differs from ATen's CPU reduction: rare 1-ulp bf16 flips allowed
    assert_bf16_close(y, v["y"], what=name)
    y2, r2 = native.add_rmsnorm(x, res, w, v["eps"])
    assert torch.equal(r2.cpu(), v["res_add"])                # the residual sum is exact
    assert_bf16_close(y2, v["y_add"], what=name + "/add")


def test_rmsnorm_layer_variant_and_strided(native):
    g = torch.Generator().manual_seed(3)
    x = torch.randn(33, 2048, generator=g).to(BF16)
    r = torch.randn(33, 2048, generator=g).to(BF16)
    w = (1 + 0.05 * torch.randn(2048, generator=g)).to(BF16)
    want, want_res = O.residual_add_then_rmsnorm(x, r, w, 1e-6)
    y, res = native.add_rmsnorm(x.to(DEV), r.to(DEV), w.to(DEV), 1e-6, round_sum_first=True)
    assert torch.equal(res.cpu(), want_res)
    assert_bf16_close(y, want, what="round_sum_first")
    wide = torch.randn(9, 4096, generator=g).to(BF16)          # a strided slice as input
    y = native.rmsnorm(wide.to(DEV)[:, 1024:3072], w.to(DEV), 1e-6)
    assert_bf16_close(y, O.rmsnorm(wide[:, 1024:3072], w, 1e-6), what="strided")
    # fused rounding: one rounding from fp32
    xf = x.float()
    fused = (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + 1e-6) * w.float()).to(BF16)
    assert_bf16_close(native.rmsnorm(x.to(DEV), w.to(DEV), 1e-6, rounding=native.ROUND_FUSED), fused, what="fused")


# ------------------------------------------------------------------------------------------ a5
@pytest.mark.parametrize("name", ["rope_1d", "mrope_interleaved", "mrope_sectioned"])
def test_rope_golden(native, golden_ops, name):
    v = golden_ops[name]
    pos = (v["positions"] if "positions" in v else v["positions3"]).to(DEV)
    mod = v["modality"].to(DEV) if "modality" in v else None
    q, k = native.qknorm_rope_kvwrite(v["q"].to(DEV), v["k"].to(DEV), None, None, None, 0.0, pos,
                                      v["inv_freq"].to(DEV), modality=mod)
    # cosf/sinf on the device vs the host libm: a last-ulp fp32 difference can flip a bf16 cos/sin
    assert_bf16_close(q, v["q_out"], max_ulp=2, max_frac=5e-3, what=name + "/q")
    assert_bf16_close(k, v["k_out"], max_ulp=2, max_frac=5e-3, what=name + "/k")


def test_qknorm_rope_kvwrite_fused(native):
    """The fused kernel == oracle q_norm/k_norm + rope + store_kvcache on the same bf16 inputs
    (models/qwen3_tts_share.py:96-106), reading q/k/v as slices of one fused QKV buffer."""
    g = torch.Generator().manual_seed(11)
    n, h, kvh, hd, page, blocks = 37, 16, 8, 128, 256, 3
    qkv = (torch.randn(n, (h + 2 * kvh) * hd, generator=g) * 1.3).to(BF16)
    qw = (1 + 0.05 * torch.randn(hd, generator=g)).to(BF16)
    kw = (1 + 0.05 * torch.randn(hd, generator=g)).to(BF16)
    pos = torch.randint(0, 4400, (n,), generator=g)
    inv = O.make_inv_freq(hd, 1e6)
    slots = torch.randperm(blocks * page, generator=g)[:n].to(torch.int32)
    slots[3] = -1
    q = qkv[:, :h * hd].view(n, h, hd)
    k = qkv[:, h * hd:(h + kvh) * hd].view(n, kvh, hd)
    v = qkv[:, (h + kvh) * hd:].view(n, kvh, hd)
    q_ref, k_ref = O.qknorm_rope(q, k, qw, kw, 1e-6, pos, inv)
    kc_ref = torch.zeros(blocks, page, kvh, hd, dtype=BF16)
    vc_ref = torch.zeros_like(kc_ref)
    O.store_kvcache(k_ref, v, kc_ref, vc_ref, slots)

    qkv_d = qkv.to(DEV)
    qd = qkv_d[:, :h * hd].view(n, h, hd)
    kd = qkv_d[:, h * hd:(h + kvh) * hd].view(n, kvh, hd)
    vd = qkv_d[:, (h + kvh) * hd:].view(n, kvh, hd)
    kc, vc = torch.zeros_like(kc_ref, device=DEV), torch.zeros_like(vc_ref, device=DEV)
    q_out, k_out = native.qknorm_rope_kvwrite(qd, kd, vd, qw.to(DEV), kw.to(DEV), 1e-6, pos.to(DEV), inv.to(DEV),
                                              slot_mapping=slots.to(DEV), k_cache=kc, v_cache=vc)
    assert_bf16_close(q_out, q_ref, max_ulp=2, max_frac=5e-3, what="q")
    assert_bf16_close(k_out, k_ref, max_ulp=2, max_frac=5e-3, what="k")
    assert torch.equal(vc.cpu(), vc_ref)                       # V is a bit copy
    # K in the cache is exactly the K the kernel returned (bit copy of its own result)
    kc_from_out = torch.zeros_like(kc_ref)
    O.store_kvcache(k_out.cpu(), v, kc_from_out, torch.zeros_like(vc_ref), slots)
    assert torch.equal(kc.cpu(), kc_from_out)


# ------------------------------------------------------------------------------------------ a6
def test_silu_golden(native, golden_ops):
    v = golden_ops["silu"]
    g, u = v["gate"].to(DEV), v["up"].to(DEV)
    assert_bf16_close(native.silu_and_mul(g, None), v["silu"], what="silu")
    assert_bf16_close(native.silu_and_mul(g, u), v["silu_mul_eager"], what="silu*mul eager")
    assert_bf16_close(native.silu_and_mul(g, u, rounding=native.ROUND_FUSED), v["silu_mul_fused"], what="fused")
    gu = torch.cat([v["gate"], v["up"]], dim=1).to(DEV)       # halves of a fused gate_up buffer
    i = v["gate"].shape[1]
    assert_bf16_close(native.silu_and_mul(gu[:, :i], gu[:, i:]), v["silu_mul_eager"], what="strided")


# ------------------------------------------------------------------------------------------ a8
@pytest.mark.parametrize("name", ["sampler_talker", "sampler_predictor", "sampler_greedy"])
def test_sampler_golden_bit_exact(native, golden_ops, name):
    v = golden_ops[name]
    probs = torch.empty_like(v["probs"], device=DEV)
    idx = native.topk_sample(v["logits"].to(DEV), v["temperatures"].to(DEV), v["noise"].to(DEV), v["top_k"],
                             probs_out=probs)
    assert torch.equal(idx.cpu(), v["index"]), (idx.cpu(), v["index"])
    # identical support (the `logits < kth` tie rule) and near-identical probabilities
    assert torch.equal(probs.cpu() > 0, v["probs"] > 0)
    assert torch.allclose(probs.cpu(), v["probs"], rtol=1e-5, atol=1e-8)


def test_sampler_random_rows_and_ties(native):
    g = torch.Generator().manual_seed(5)
    b, vsz = 97, 3072
    logits = (torch.randn(b, vsz, generator=g) * 2).to(BF16)
    logits[:, 100:140] = logits[:, 99:100]                    # 41-way tie crossing the k-th place
    temps = torch.rand(b, generator=g) * 1.5 + 0.2
    noise = torch.empty(b, vsz).exponential_(1, generator=g)
    for top_k in (1, 50, 4000):
        want = O.topk_sample(logits, temps, noise, min(top_k, vsz))
        got = native.topk_sample(logits.to(DEV), temps.to(DEV), noise.to(DEV), top_k)
        assert torch.equal(got.cpu(), want), top_k
    got32 = native.topk_sample(logits.float().to(DEV), temps.to(DEV), noise.to(DEV), 50)
    assert torch.equal(got32.cpu(), O.topk_sample(logits, temps, noise, 50))


@pytest.mark.parametrize("b,vsz,top_k", [(64, 2048, 50), (3, 3072, 50), (5, 4096, 7), (4, 1000, 50), (2, 5000, 50),
                                        (7, 512, 600), (1, 2048, 1)])
def test_sampler_warp_and_block_paths_with_fused_embedding(native, b, vsz, top_k):
    """The warp-per-row kernel (vocab <= 4096, multiple of 8) and the block kernel (everything else) against the
    oracle, fp32 and bf16 logits, strided logits rows, heavy ties; and the fused gather of the chosen code's
    embedding row (models/qwen3_tts_predictor.py:94-95)."""
    g = torch.Generator().manual_seed(b * 31 + vsz)
    logits = (torch.randn(b, vsz, generator=g) * 3).to(BF16)
    logits[:, 10:60] = logits[:, 9:10]                           # 51-way tie, crosses the k-th place for k = 50
    temps = torch.rand(b, generator=g) * 1.5 + 0.2
    noise = torch.empty(b, vsz).exponential_(1, generator=g)
    want = O.topk_sample(logits, temps, noise, min(top_k, vsz))
    got = native.topk_sample(logits.to(DEV), temps.to(DEV), noise.to(DEV), top_k)
    assert torch.equal(got.cpu(), want)
    wide = torch.zeros(b, vsz + 64, dtype=BF16, device=DEV)       # rows of a wider buffer (row stride != vocab)
    wide[:, :vsz] = logits.to(DEV)
    assert torch.equal(native.topk_sample(wide[:, :vsz], temps.to(DEV), noise.to(DEV), top_k).cpu(), want)
    got32 = native.topk_sample(logits.float().to(DEV), temps.to(DEV), noise.to(DEV), top_k)
    assert torch.equal(got32.cpu(), want)
    if vsz <= 4096 and vsz % 8 == 0:
        table = torch.randn(vsz, 256, generator=g).to(BF16).to(DEV)
        dst = torch.zeros(b, 2, 256, dtype=BF16, device=DEV)
        idx = native.topk_sample(logits.to(DEV), temps.to(DEV), noise.to(DEV), top_k, embed_table=table, embed_out=dst[:, 1])
        assert torch.equal(idx.cpu(), want)
        assert torch.equal(dst[:, 1], table[idx]) and dst[:, 0].abs().max().item() == 0


def test_sampler_bf16_select_negative_thresholds_and_padding(native):
    """bf16 logits go through a 16-bit radix select in logit space (k-th largest x, then x / T): thresholds that are
    negative, rows of all-equal values, -inf entries, k close to the vocabulary and a vocabulary that does not fill the
    last chunk of a thread block."""
    g = torch.Generator().manual_seed(77)
    for vsz, top_k in ((2048, 50), (2048, 2000), (1000, 50), (3072, 3071), (4096, 2)):
        b = 9
        logits = (torch.randn(b, vsz, generator=g) * 2 - 30).to(BF16)          # every threshold negative
        logits[1] = -7.25                                                       # one value everywhere
        logits[2, ::3] = float("-inf")
        logits[3] = (torch.randn(vsz, generator=g) * 1e-3).to(BF16)            # thresholds around +-0
        logits[4, : vsz // 2] = 0.0
        logits[4, vsz // 2:] = -0.0
        temps = torch.rand(b, generator=g) * 3 + 0.05
        noise = torch.empty(b, vsz).exponential_(1, generator=g)
        want = O.topk_sample(logits, temps, noise, min(top_k, vsz))
        got = native.topk_sample(logits.to(DEV), temps.to(DEV), noise.to(DEV), top_k)
        assert torch.equal(got.cpu(), want), (vsz, top_k)
        got32 = native.topk_sample(logits.float().to(DEV), temps.to(DEV), noise.to(DEV), top_k)
        assert torch.equal(got32.cpu(), want), (vsz, top_k)


# ------------------------------------------------------------------------------------------ a2
def _paged_case(b, lens, page=256, kvh=8, h=16, hd=128, seed=0, extra_pages=2):
    g = torch.Generator().manual_seed(seed)
    pages_per = [(l + page - 1) // page for l in lens]
    nblocks = sum(pages_per) + extra_pages
    perm = torch.randperm(nblocks, generator=g).tolist()       # non-contiguous pages
    width = max(max(pages_per), 1) + 1
    bt = torch.full((b, width), -1, dtype=torch.int32)
    it = iter(perm)
    for i, n in enumerate(pages_per):
        for j in range(n):
            bt[i, j] = next(it)
    kc = torch.randn(nblocks, page, kvh, hd, generator=g).to(BF16)
    vc = torch.randn(nblocks, page, kvh, hd, generator=g).to(BF16)
    q = torch.randn(b, h, hd, generator=g).to(BF16)
    return q, kc, vc, bt, torch.tensor(lens, dtype=torch.int32)


@pytest.mark.parametrize("lens,max_ctx", [
    ([1], 256), ([3, 16, 7, 9], 17), ([64, 65, 255, 256, 257, 576], 1024), ([0, 5, 0, 300], 512),
    ([4352], 4608), ([1000] * 3, 1024),
])
def test_paged_attn_decode_vs_oracle(native, lens, max_ctx):
    q, kc, vc, bt, cl = _paged_case(len(lens), lens, seed=len(lens) + max_ctx)
    scale = 128 ** -0.5
    want = O.paged_attn_decode(q, kc, vc, cl, bt, scale)
    got = native.paged_attn_decode(q.to(DEV), kc.to(DEV), vc.to(DEV), bt.to(DEV), cl.to(DEV), scale, max_ctx)
    assert rel_err(got, want) <= 2e-2, rel_err(got, want)      # north_star tolerance
    assert rel_err(got, want) <= 8e-3                           # and in practice one bf16 ulp
    for i, l in enumerate(lens):
        if l == 0:
            assert float(got[i].abs().max()) == 0.0             # graph-padding rows give zeros


def test_paged_attn_decode_full_size_properties(native):
    """BASELINE C5 sizes (B=64 ragged contexts): size-independent properties instead of the
    oracle - (i) V == const => output == const; (ii) softmax over identical keys == mean of V;
    (iii) result independent of the split-KV factor (max_ctx changes nsplit)."""
    g = torch.Generator().manual_seed(9)
    b = 64
    lens = torch.randint(64, 1024, (b,), generator=g).tolist()
    q, kc, vc, bt, cl = _paged_case(b, lens, seed=1)
    scale = 128 ** -0.5
    qd, kd, vd, btd, cld = q.to(DEV), kc.to(DEV), vc.to(DEV), bt.to(DEV), cl.to(DEV)
    ones = torch.full_like(vd, 0.75)
    o = native.paged_attn_decode(qd, kd, ones, btd, cld, scale, 1024)
    assert torch.equal(o, torch.full_like(o, 0.75))
    same_k = torch.zeros_like(kd)
    o = native.paged_attn_decode(qd, same_k, vd, btd, cld, scale, 1024)
    for i in (0, 17, 63):
        pages = bt[i, :(lens[i] + 255) // 256].long()
        mean_v = vc[pages].reshape(-1, 8, 128)[:lens[i]].float().mean(0).repeat_interleave(2, 0)
        assert rel_err(o[i], mean_v) <= 2e-2
    a = native.paged_attn_decode(qd, kd, vd, btd, cld, scale, 1024)
    c = native.paged_attn_decode(qd, kd, vd, btd, cld, scale, 65536)
    assert rel_err(a, c) <= 8e-3


def test_paged_attn_decode_vs_flash_attn(native):
    """Cross-check against the kernel the reference itself calls (attention.py:93-95)."""
    fa = pytest.importorskip("flash_attn")
    lens = [64, 300, 576, 1, 1024, 77, 256, 900]
    q, kc, vc, bt, cl = _paged_case(len(lens), lens, seed=4)
    scale = 128 ** -0.5
    qd, kd, vd, btd, cld = q.to(DEV), kc.to(DEV), vc.to(DEV), bt.clamp_min(0).to(DEV), cl.to(DEV)
    try:
        want = fa.flash_attn_with_kvcache(qd.unsqueeze(1), kd, vd, cache_seqlens=cld, block_table=btd,
                                          softmax_scale=scale, causal=True).squeeze(1)
    except Exception as e:                                       # wheel without a kernel for this GPU
        pytest.skip(f"flash_attn unavailable here: {e}")
    got = native.paged_attn_decode(qd, kd, vd, btd, cld, scale, 1024)
    assert rel_err(got, want) <= 2e-2


# ------------------------------------------------------------------------------------------ a3
@pytest.mark.parametrize("lens", [[2, 2, 2], [11, 9], [1], [40], [8] * 16, [300, 5]])
def test_varlen_prefill_vs_oracle(native, lens):
    g = torch.Generator().manual_seed(sum(lens))
    t, h, kvh, hd = sum(lens), 16, 8, 128
    qkv = torch.randn(t, (h + 2 * kvh) * hd, generator=g).to(BF16)
    cu = torch.tensor([0] + torch.tensor(lens).cumsum(0).tolist(), dtype=torch.int32)
    q = qkv[:, :h * hd].view(t, h, hd)
    k = qkv[:, h * hd:(h + kvh) * hd].view(t, kvh, hd)
    v = qkv[:, (h + kvh) * hd:].view(t, kvh, hd)
    scale = hd ** -0.5
    want = O.varlen_attn_prefill(q, k, v, cu, cu, scale)
    d = qkv.to(DEV)
    qd, kd, vd = d[:, :h * hd].view(t, h, hd), d[:, h * hd:(h + kvh) * hd].view(t, kvh, hd), d[:, (h + kvh) * hd:].view(t, kvh, hd)
    got = native.varlen_attn_prefill(qd, kd, vd, cu.to(DEV), cu.to(DEV), max(lens), max(lens), scale)   # strided rows
    assert rel_err(got, want) <= 8e-3
    got = native.varlen_attn_prefill(qd.contiguous(), kd.contiguous(), vd.contiguous(), cu.to(DEV), cu.to(DEV),
                                     max(lens), max(lens), scale)                                         # contiguous rows
    assert rel_err(got, want) <= 8e-3


def test_varlen_prefill_paged_prefix(native):
    """block_table branch (attention.py:79-80): queries are the tail of a longer cached context."""
    lens_k, lens_q = [300, 64], [44, 64]
    q_, kc, vc, bt, _ = _paged_case(2, lens_k, seed=21)
    g = torch.Generator().manual_seed(2)
    q = torch.randn(sum(lens_q), 16, 128, generator=g).to(BF16)
    cuq = torch.tensor([0, 44, 108], dtype=torch.int32)
    cuk = torch.tensor([0, 300, 364], dtype=torch.int32)
    scale = 128 ** -0.5
    want = O.varlen_attn_prefill(q, None, None, cuq, cuk, scale, kc, vc, bt)
    got = native.varlen_attn_prefill(q.to(DEV), kc.to(DEV), vc.to(DEV), cuq.to(DEV), cuk.to(DEV), 64, 300, scale,
                                     block_table=bt.to(DEV))
    assert rel_err(got, want) <= 8e-3


# ------------------------------------------------------------------------------------------ a9
def test_gathers(native):
    g = torch.Generator().manual_seed(8)
    tabs = torch.randn(15, 2048, 256, generator=g).to(BF16)
    rows = torch.randint(0, 2048, (23,), generator=g)
    tix = torch.randint(0, 15, (23,), generator=g).to(torch.int32)
    got = native.gather_rows(tabs.to(DEV), rows.to(DEV), tix.to(DEV))
    assert torch.equal(got.cpu(), tabs[tix.long(), rows])
    t0 = torch.randn(3072, 256, generator=g).to(BF16)
    codes = torch.cat([torch.randint(0, 3072, (23, 1), generator=g), torch.randint(0, 2048, (23, 15), generator=g)], 1)
    extra = torch.randn(1, 256, generator=g).to(BF16)
    want = torch.stack([torch.cat([t0[codes[i, 0]][None], tabs[torch.arange(15), codes[i, 1:]]], 0).unsqueeze(0).sum(1)[0]
                        for i in range(23)]) + extra
    got = native.codec_embed_sum(t0.to(DEV), tabs.to(DEV), codes.to(DEV), extra.to(DEV))
    assert_bf16_close(got, want, what="codec_embed_sum")
