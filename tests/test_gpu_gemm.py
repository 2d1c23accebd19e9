"""GPU parity of the weight-streaming tcgen05 GEMM (b2tts_gemm_bf16) through the C ABI.

Reference semantics: ``F.linear`` (layers/linear.py:72-73,149-153) in bf16 with fp32 accumulation; the
fused epilogues restate ``residual + hidden`` (models/qwen3_tts_share.py:171,175) and
``act_fn(gate) * up`` (models/qwen3_tts_share.py:131-135).  Tolerance: logits/projections within 2e-2
relative (north_star); in practice the only difference from an fp64 contraction is the fp32 summation
order, so the bound used here is one bf16 ulp on almost every element."""
import pytest
import torch

pytestmark = pytest.mark.gpu
BF16 = torch.bfloat16


def _ref_linear(x, w, bias=None):
    y = x.double() @ w.double().t()
    if bias is not None:
        y = y + bias.double()
    return y


def _check(got, want64, what, max_rel=2e-2):
    want = want64.to(BF16).float()
    err = (got.float() - want64.float()).abs().max().item()
    scale = want64.abs().max().item() + 1e-12
    assert err / scale <= max_rel, f"{what}: max rel err {err / scale:.3e}"
    # bf16 rounding of an fp32 sum: at most ~1 ulp away from the rounded fp64 result
    ulp = (got.float() - want).abs() / (want.abs().clamp_min(1e-3) * 2 ** -8)
    frac_bad = (ulp > 1.01).float().mean().item()
    assert frac_bad < 2e-3, f"{what}: {frac_bad:.4f} of elements differ by more than one bf16 ulp"


SHAPES = [  # (n, k): every projection of the 1.7B talker and the code predictor
    (4096, 2048), (2048, 2048), (2048, 6144), (3072, 2048),
    (1024, 2048), (4096, 1024), (1024, 3072), (2048, 1024),
]


@pytest.mark.parametrize("n,k", SHAPES)
@pytest.mark.parametrize("m", [1, 8, 64])
def test_gemm_plain(native, m, n, k):
    g = torch.Generator(device="cuda").manual_seed(m * 7 + n + k)
    x = torch.randn(m, k, device="cuda", generator=g).to(BF16)
    w = (torch.randn(n, k, device="cuda", generator=g) * 0.05).to(BF16)
    packed = native.gemm_pack_weight(w)
    y = native.gemm_bf16(x, packed, n)
    _check(y, _ref_linear(x, w), f"gemm m={m} n={n} k={k}")


@pytest.mark.parametrize("m", [3, 17, 100, 128])
def test_gemm_ragged_m_and_bias(native, m):
    n, k = 1024, 2048
    g = torch.Generator(device="cuda").manual_seed(m)
    x = torch.randn(m, k, device="cuda", generator=g).to(BF16)
    w = (torch.randn(n, k, device="cuda", generator=g) * 0.05).to(BF16)
    b = torch.randn(n, device="cuda", generator=g).to(BF16)
    packed = native.gemm_pack_weight(w)
    y = native.gemm_bf16(x, packed, n, bias=b)
    _check(y, _ref_linear(x, w, b), f"gemm+bias m={m}")


@pytest.mark.parametrize("splits", [1, 2, 4, 8, 16])
def test_gemm_every_cluster_size_is_deterministic(native, splits):
    m, n, k = 64, 1024, 2048
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn(m, k, device="cuda", generator=g).to(BF16)
    w = (torch.randn(n, k, device="cuda", generator=g) * 0.05).to(BF16)
    packed = native.gemm_pack_weight(w)
    y1 = native.gemm_bf16(x, packed, n, splits=splits)
    y2 = native.gemm_bf16(x, packed, n, splits=splits)
    assert torch.equal(y1, y2)
    _check(y1, _ref_linear(x, w), f"gemm splits={splits}")


def test_gemm_strided_input_and_pdl_flag(native):
    m, n, k = 32, 2048, 1024
    g = torch.Generator(device="cuda").manual_seed(9)
    big = torch.randn(m, k + 64, device="cuda", generator=g).to(BF16)
    x = big[:, 64:]                                            # row stride k+64, 16-byte aligned
    w = (torch.randn(n, k, device="cuda", generator=g) * 0.05).to(BF16)
    packed = native.gemm_pack_weight(w)
    y = native.gemm_bf16(x, packed, n, pdl=True)
    _check(y, _ref_linear(x, w), "gemm strided + pdl")


def test_gemm_residual_epilogue_matches_reference_rounding(native):
    """bf16(residual + bf16(x W^T)) - the decoder layer's un-fused add (qwen3_tts_share.py:171,175)."""
    m, n, k = 64, 2048, 6144
    g = torch.Generator(device="cuda").manual_seed(3)
    x = torch.randn(m, k, device="cuda", generator=g).to(BF16)
    w = (torch.randn(n, k, device="cuda", generator=g) * 0.02).to(BF16)
    res = torch.randn(m, n, device="cuda", generator=g).to(BF16)
    packed = native.gemm_pack_weight(w)
    plain = native.gemm_bf16(x, packed, n)
    want = res + plain                                          # bf16 add of the rounded projection
    got = native.gemm_bf16(x, packed, n, residual=res, epilogue=native.GEMM_RESIDUAL)
    assert torch.equal(got, want)
    inplace = res.clone()
    native.gemm_bf16(x, packed, n, residual=inplace, epilogue=native.GEMM_RESIDUAL, out=inplace)
    assert torch.equal(inplace, want)


@pytest.mark.parametrize("rounding", [0, 1])
def test_gemm_silu_mul_epilogue_equals_separate_kernels(native, rounding):
    """gate_up GEMM + SiLU*mul in one launch == GEMM, then b2tts_silu_and_mul on its bf16 output."""
    m, inter, k = 64, 3072, 1024
    g = torch.Generator(device="cuda").manual_seed(4)
    x = torch.randn(m, k, device="cuda", generator=g).to(BF16)
    w = (torch.randn(2 * inter, k, device="cuda", generator=g) * 0.05).to(BF16)     # [gate; up]
    plain = native.gemm_bf16(x, native.gemm_pack_weight(w), 2 * inter)
    want = native.silu_and_mul(plain[:, :inter], plain[:, inter:], rounding=rounding)
    got = native.gemm_bf16(x, native.gemm_pack_weight(w, native.PACK_GATE_UP), 2 * inter,
                           epilogue=native.GEMM_SILU_MUL, rounding=rounding)
    assert got.shape == (m, inter)
    assert torch.equal(got, want)


def test_gemm_linearity_at_full_size(native):
    """Size-independent property at the largest projection (gate_up of the 1.7B talker, M=128):
    the K-split partials add up - y(x1) + y(x2) == y(x1 + x2) when every product is exact."""
    m, n, k = 128, 12288, 2048
    g = torch.Generator(device="cuda").manual_seed(8)
    # small integers: every product and partial sum is exact in fp32 and in bf16 output range
    x1 = torch.randint(-2, 3, (m, k), device="cuda", generator=g).to(BF16)
    x2 = torch.randint(-2, 3, (m, k), device="cuda", generator=g).to(BF16)
    w = torch.randint(-1, 2, (n, k), device="cuda", generator=g).to(BF16)
    w[:, ::2] = 0                                              # keep |y| < 256 * ... exactly representable
    w[:, 1::4] = 0
    packed = native.gemm_pack_weight(w)
    y1 = native.gemm_bf16(x1, packed, n).float()
    y2 = native.gemm_bf16(x2, packed, n).float()
    y12 = native.gemm_bf16(x1 + x2, packed, n).float()
    want = (x1.float() @ w.float().t())
    assert want.abs().max() < 256 * 8
    exact = want.to(BF16).float() == want
    assert torch.equal(y1[exact], want[exact])
    both = exact & ((y1 + y2).to(BF16).float() == (y1 + y2))
    assert torch.equal(y12[both], (y1 + y2)[both])


def test_gemm_argument_errors(native):
    x = torch.zeros(4, 100, device="cuda", dtype=BF16)
    assert not native.gemm_supported(4, 128, 100)
    assert native.gemm_supported(300, 128, 64)              # rows beyond 128 run as 128-row chunks of the grid
    assert not native.gemm_supported((1 << 16) + 1, 128, 64)
    with pytest.raises(RuntimeError):
        native.gemm_pack_weight(torch.zeros(100, 64, device="cuda", dtype=BF16))
    w = native.gemm_pack_weight(torch.zeros(128, 128, device="cuda", dtype=BF16))
    with pytest.raises(RuntimeError):
        native.gemm_bf16(x, w, 128)


# ------------------------------------------------------------------ fused RMSNorm in, row sums of squares out
@pytest.mark.parametrize("m", [1, 8, 64, 100, 128])
@pytest.mark.parametrize("rounding", [0, 1])
def test_gemm_ssq_out_then_fused_norm_in(native, m, rounding):
    """down/o-style GEMM with the residual epilogue emits per-row partial sums of squares; the next
    GEMM normalises its input with them while staging.  Must equal rmsnorm kernel + plain GEMM
    (same rounding points) up to the fp32 order of the row sum."""
    h, inter, eps = 2048, 3072, 1e-6
    g = torch.Generator(device="cuda").manual_seed(m + 31 * rounding)
    act = torch.randn(m, inter, device="cuda", generator=g).to(BF16)
    res = torch.randn(m, h, device="cuda", generator=g).to(BF16)
    w_down = (torch.randn(h, inter, device="cuda", generator=g) * 0.03).to(BF16)
    w_qkv = (torch.randn(4096, h, device="cuda", generator=g) * 0.03).to(BF16)
    nw = (1 + 0.1 * torch.randn(h, device="cuda", generator=g)).to(BF16)
    pd, pq = native.gemm_pack_weight(w_down), native.gemm_pack_weight(w_qkv)
    slots = native.gemm_ssq_slots(m, h, inter)
    ssq = torch.full((m, slots), float("nan"), device="cuda")
    new_res = native.gemm_bf16(act, pd, h, residual=res, epilogue=native.GEMM_RESIDUAL, ssq_out=ssq)
    want_res = native.gemm_bf16(act, pd, h, residual=res, epilogue=native.GEMM_RESIDUAL)
    assert torch.equal(new_res, want_res)
    torch.testing.assert_close(ssq.sum(1), new_res.float().pow(2).sum(1), rtol=1e-5, atol=1e-5)
    y = native.gemm_bf16(new_res, pq, 4096, rounding=rounding, norm_weight=nw, ssq_in=ssq, eps=eps)
    normed = native.rmsnorm(new_res, nw, eps, rounding)
    want = native.gemm_bf16(normed, pq, 4096)
    # identical unless the differently-ordered row sum moves rstd by an ulp and flips a bf16 rounding
    diff = (y.float() - want.float()).abs()
    assert diff.max().item() <= 2e-2 * want.float().abs().max().item()
    assert (diff > 0).float().mean().item() < 0.02
    y2 = native.gemm_bf16(new_res, pq, 4096, rounding=rounding, norm_weight=nw, ssq_in=ssq, eps=eps)
    assert torch.equal(y, y2)


@pytest.mark.parametrize("slots", [1, 4, 8, 2])
def test_gemm_fused_norm_any_slot_count(native, slots):
    m, h, eps = 24, 1024, 1e-6
    g = torch.Generator(device="cuda").manual_seed(slots)
    x = torch.randn(m, h, device="cuda", generator=g).to(BF16)
    nw = (1 + 0.1 * torch.randn(h, device="cuda", generator=g)).to(BF16)
    w = (torch.randn(3072 * 2, h, device="cuda", generator=g) * 0.03).to(BF16)
    ssq = x.float().pow(2).view(m, slots, -1).sum(2).contiguous()
    pk = native.gemm_pack_weight(w, native.PACK_GATE_UP)
    y = native.gemm_bf16(x, pk, 6144, epilogue=native.GEMM_SILU_MUL, norm_weight=nw, ssq_in=ssq, eps=eps)
    normed = native.rmsnorm(x, nw, eps)
    want = native.gemm_bf16(normed, pk, 6144, epilogue=native.GEMM_SILU_MUL)
    diff = (y.float() - want.float()).abs()
    assert diff.max().item() <= 2e-2 * want.float().abs().max().item()
    assert (diff > 0).float().mean().item() < 0.02


def test_gemm_bias_epilogue_ssq_out(native):
    m, n, k = 64, 1024, 2048
    g = torch.Generator(device="cuda").manual_seed(9)
    x = torch.randn(m, k, device="cuda", generator=g).to(BF16)
    w = (torch.randn(n, k, device="cuda", generator=g) * 0.05).to(BF16)
    b = torch.randn(n, device="cuda", generator=g).to(BF16)
    ssq = torch.zeros(m, native.gemm_ssq_slots(m, n, k), device="cuda")
    y = native.gemm_bf16(x, native.gemm_pack_weight(w), n, bias=b, ssq_out=ssq)
    assert torch.equal(y, native.gemm_bf16(x, native.gemm_pack_weight(w), n, bias=b))
    torch.testing.assert_close(ssq.sum(1), y.float().pow(2).sum(1), rtol=1e-5, atol=1e-5)


# ------------------------------------------------------------------ more than 128 rows: grid chunks of 128 rows
@pytest.mark.parametrize("m", [129, 256, 300, 1000, 4096])
def test_gemm_row_chunks(native, m):
    """Prefill-sized M (SURVEY.md section 8a row a7 "lift G_MAX_M"): every 128-row chunk is its own set of CTAs,
    the last one ragged.  Same bound as the single-chunk case, every epilogue."""
    n, k = 2048, 1024
    g = torch.Generator(device="cuda").manual_seed(m)
    x = torch.randn(m, k, device="cuda", generator=g).to(BF16)
    w = (torch.randn(n, k, device="cuda", generator=g) * 0.05).to(BF16)
    b = torch.randn(n, device="cuda", generator=g).to(BF16)
    res = torch.randn(m, n, device="cuda", generator=g).to(BF16)
    pk = native.gemm_pack_weight(w)
    _check(native.gemm_bf16(x, pk, n, bias=b), _ref_linear(x, w, b), f"chunked gemm+bias m={m}")
    y = native.gemm_bf16(x, pk, n, residual=res, epilogue=native.GEMM_RESIDUAL)
    want = res.float() + (x.double() @ w.double().t()).to(BF16).float()
    assert (y.float() - want).abs().max().item() <= 2e-2 * want.abs().max().item()
    # rows are independent: chunked result == the same rows computed 128 at a time (same K split => same fp32 order)
    full = native.gemm_bf16(x, pk, n, bias=b, splits=2)
    for lo in (0, 128 * ((m - 1) // 128)):
        hi = min(m, lo + 128)
        assert torch.equal(native.gemm_bf16(x[lo:hi], pk, n, bias=b, splits=2), full[lo:hi])


@pytest.mark.parametrize("m", [200, 512])
@pytest.mark.parametrize("rounding", [0, 1])
def test_gemm_row_chunks_fused_norm_ssq_silu(native, m, rounding):
    h, inter, eps = 1024, 3072, 1e-6
    g = torch.Generator(device="cuda").manual_seed(m + rounding)
    act = torch.randn(m, inter, device="cuda", generator=g).to(BF16)
    res = torch.randn(m, h, device="cuda", generator=g).to(BF16)
    w_down = (torch.randn(h, inter, device="cuda", generator=g) * 0.03).to(BF16)
    w_gu = (torch.randn(2 * inter, h, device="cuda", generator=g) * 0.03).to(BF16)
    nw = (1 + 0.1 * torch.randn(h, device="cuda", generator=g)).to(BF16)
    pd, pg = native.gemm_pack_weight(w_down), native.gemm_pack_weight(w_gu, native.PACK_GATE_UP)
    ssq = torch.full((m, native.gemm_ssq_slots(m, h, inter)), float("nan"), device="cuda")
    new_res = native.gemm_bf16(act, pd, h, residual=res, epilogue=native.GEMM_RESIDUAL, ssq_out=ssq)
    assert torch.equal(new_res, native.gemm_bf16(act, pd, h, residual=res, epilogue=native.GEMM_RESIDUAL))
    torch.testing.assert_close(ssq.sum(1), new_res.float().pow(2).sum(1), rtol=1e-5, atol=1e-5)
    y = native.gemm_bf16(new_res, pg, 2 * inter, epilogue=native.GEMM_SILU_MUL, rounding=rounding, norm_weight=nw,
                         ssq_in=ssq, eps=eps)
    want = native.gemm_bf16(native.rmsnorm(new_res, nw, eps, rounding), pg, 2 * inter, epilogue=native.GEMM_SILU_MUL,
                            rounding=rounding)
    assert y.shape == (m, inter)
    diff = (y.float() - want.float()).abs()
    assert diff.max().item() <= 2e-2 * want.float().abs().max().item()
    assert (diff > 0).float().mean().item() < 0.02


def test_gemm_prefetch_next_weights_changes_nothing(native):
    """b2tts_gemm_prefetch_bf16: the L2 prefetch of the next projection's packed weights is a hint only."""
    m, n, k = 64, 1024, 2048
    g = torch.Generator(device="cuda").manual_seed(77)
    x = torch.randn(m, k, device="cuda", generator=g).to(BF16)
    w = (torch.randn(n, k, device="cuda", generator=g) * 0.05).to(BF16)
    nxt = native.gemm_pack_weight((torch.randn(4096, n, device="cuda", generator=g) * 0.05).to(BF16))
    pk = native.gemm_pack_weight(w)
    ssq = torch.zeros(m, native.gemm_ssq_slots(m, n, k), device="cuda")
    y = native.gemm_bf16(x, pk, n, ssq_out=ssq, prefetch=nxt)
    assert torch.equal(y, native.gemm_bf16(x, pk, n))
    assert torch.equal(native.gemm_bf16(x, pk, n, prefetch=nxt), y)
    with pytest.raises(RuntimeError):
        native.gemm_bf16(x, pk, n, prefetch=nxt.view(-1)[8:8 + 16384])        # not 16-byte aligned


# ------------------------------------------------------------------ batch split (splits < 0): W multicast, no K split
@pytest.mark.parametrize("m,bs", [(64, 4), (50, 4), (32, 2), (17, 2), (128, 4), (96, 2), (64, 2)])
@pytest.mark.parametrize("n,k", [(4096, 1024), (2048, 1024), (1024, 2048), (256, 128), (128, 64)])
def test_gemm_batch_split_plain_bias_residual(native, m, bs, n, k):
    """The cluster splits the batch (each CTA: 16 - 48 rows x the whole K, weight tiles multicast): one fp32 accumulator
    over K in order, so against fp64 it obeys the same bound as the K-split kernel; run-to-run identical."""
    g = torch.Generator(device="cuda").manual_seed(m * 13 + n + k + bs)
    x = torch.randn(m, k, device="cuda", generator=g).to(BF16)
    w = (torch.randn(n, k, device="cuda", generator=g) * 0.05).to(BF16)
    b = torch.randn(n, device="cuda", generator=g).to(BF16)
    res = torch.randn(m, n, device="cuda", generator=g).to(BF16)
    packed = native.gemm_pack_weight(w)
    y = native.gemm_bf16(x, packed, n, splits=-bs)
    _check(y, _ref_linear(x, w), f"batch-split gemm m={m} n={n} k={k} bs={bs}")
    assert torch.equal(y, native.gemm_bf16(x, packed, n, splits=-bs, pdl=True))
    _check(native.gemm_bf16(x, packed, n, bias=b, splits=-bs), _ref_linear(x, w, b), "batch-split gemm + bias")
    # residual epilogue: bf16(res + bf16(acc)); identical to the K-split kernel's wherever bf16(acc) is
    yr = native.gemm_bf16(x, packed, n, residual=res, epilogue=native.GEMM_RESIDUAL, splits=-bs)
    want = (res.float() + _ref_linear(x, w).to(BF16).float()).to(BF16)
    assert ((yr.float() - want.float()).abs() > 0).float().mean().item() < 5e-3
    ssq = torch.full((m, native.gemm_ssq_slots(m, n, k)), float("nan"), device="cuda")
    yr2 = native.gemm_bf16(x, packed, n, residual=res, epilogue=native.GEMM_RESIDUAL, splits=-bs, ssq_out=ssq)
    assert torch.equal(yr, yr2)
    torch.testing.assert_close(ssq.sum(1), yr.float().pow(2).sum(1), rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("m,bs", [(64, 4), (40 + 24, 4), (32, 2), (128, 4), (20, 2)])
@pytest.mark.parametrize("rounding", [0, 1])
def test_gemm_batch_split_fused_norm_and_silu_mul(native, m, bs, rounding):
    """Fused input RMSNorm (ssq_in) and the SiLU*mul epilogue in batch-split mode against the separate kernels."""
    h, inter, eps = 1024, 3072, 1e-6
    g = torch.Generator(device="cuda").manual_seed(m + 7 * rounding + bs)
    x = torch.randn(m, h, device="cuda", generator=g).to(BF16)
    nw = (1 + 0.1 * torch.randn(h, device="cuda", generator=g)).to(BF16)
    w = (torch.randn(2 * inter, h, device="cuda", generator=g) * 0.03).to(BF16)
    wq = (torch.randn(4096, h, device="cuda", generator=g) * 0.03).to(BF16)
    pk, pq = native.gemm_pack_weight(w, native.PACK_GATE_UP), native.gemm_pack_weight(wq)
    ssq = x.float().pow(2).view(m, 8, -1).sum(2).contiguous()
    normed = native.rmsnorm(x, nw, eps, rounding)
    for packed, n, epi in ((pk, 2 * inter, native.GEMM_SILU_MUL), (pq, 4096, native.GEMM_BIAS)):
        y = native.gemm_bf16(x, packed, n, epilogue=epi, rounding=rounding, norm_weight=nw, ssq_in=ssq, eps=eps, splits=-bs)
        want = native.gemm_bf16(normed, packed, n, epilogue=epi, rounding=rounding, splits=-bs)
        diff = (y.float() - want.float()).abs()
        assert diff.max().item() <= 2e-2 * want.float().abs().max().item()
        assert (diff > 0).float().mean().item() < 0.02
        ksplit = native.gemm_bf16(normed, packed, n, epilogue=epi, rounding=rounding, splits=4)
        d2 = (want.float() - ksplit.float()).abs()
        assert d2.max().item() <= 2e-2 * ksplit.float().abs().max().item()
        assert torch.equal(y, native.gemm_bf16(x, packed, n, epilogue=epi, rounding=rounding, norm_weight=nw, ssq_in=ssq,
                                               eps=eps, splits=-bs, pdl=True))


def test_gemm_batch_split_argument_errors(native):
    x = torch.zeros(40, 1024, device="cuda", dtype=BF16)
    w = native.gemm_pack_weight(torch.zeros(128, 1024, device="cuda", dtype=BF16))
    for bad in (-4, -2, -3, -8):            # 40 rows pad to 48: not a multiple of 32 / 64; 3 and 8 are not cluster sizes
        with pytest.raises(RuntimeError):
            native.gemm_bf16(x, w, 128, splits=bad)
