"""GPU parity of the persistent GEMM chain (b2tts_gemm_chain, SURVEY.md section 8a rows a7 + a4 + a6) through
the C ABI: the four dependent projections of a decoder layer in one launch must give what the four separate
b2tts_gemm_norm_bf16 launches give (same packed weights, epilogues, fused norms and rounding points; the only
difference allowed is the fp32 summation order of the K splits: one bf16 ulp on a few elements of each op's
output, compounding along the chain, bounded here at 1e-2 of the output's magnitude - inside the north_star
tolerance of 2e-2; the separate launches are themselves checked against fp64 in tests/test_gpu_gemm.py)."""
import pytest
import torch

pytestmark = pytest.mark.gpu
BF16 = torch.bfloat16
EPS = 1e-6


def _layer_weights(h, inter, qkv, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    mk = lambda n, k: (torch.randn(n, k, device="cuda", generator=g) * 0.03).to(BF16)
    return dict(o=mk(h, qkv // 2), gu=mk(2 * inter, h), down=mk(h, inter), qkv=mk(qkv, h),
                ln1=(1 + 0.1 * torch.randn(h, device="cuda", generator=g)).to(BF16),
                ln2=(1 + 0.1 * torch.randn(h, device="cuda", generator=g)).to(BF16))


def _separate(native, w, attn, resid, rounding):
    """The chain's four ops as four b2tts_gemm_norm_bf16 launches."""
    m = attn.shape[0]
    h = w["o"].shape[0]
    po, pgu, pd, pq = (native.gemm_pack_weight(w["o"]), native.gemm_pack_weight(w["gu"], native.PACK_GATE_UP),
                       native.gemm_pack_weight(w["down"]), native.gemm_pack_weight(w["qkv"]))
    ssq1 = torch.empty(m, native.gemm_ssq_slots(m, h, attn.shape[1]), dtype=torch.float32, device="cuda")
    r1 = native.gemm_bf16(attn, po, h, residual=resid, epilogue=native.GEMM_RESIDUAL, ssq_out=ssq1)
    act = native.gemm_bf16(r1, pgu, w["gu"].shape[0], epilogue=native.GEMM_SILU_MUL, rounding=rounding,
                           norm_weight=w["ln1"], ssq_in=ssq1, eps=EPS)
    ssq2 = torch.empty(m, native.gemm_ssq_slots(m, h, act.shape[1]), dtype=torch.float32, device="cuda")
    r2 = native.gemm_bf16(act, pd, h, residual=r1, epilogue=native.GEMM_RESIDUAL, ssq_out=ssq2)
    qkv = native.gemm_bf16(r2, pq, w["qkv"].shape[0], norm_weight=w["ln2"], ssq_in=ssq2, eps=EPS, rounding=rounding)
    return r1, act, r2, qkv


def _chain(native, w, attn, resid, rounding, nops=4, in_place=False):
    m = attn.shape[0]
    h, inter, nq = w["o"].shape[0], w["down"].shape[1], w["qkv"].shape[0]
    packed = dict(o=native.gemm_pack_weight(w["o"]), gu=native.gemm_pack_weight(w["gu"], native.PACK_GATE_UP),
                  down=native.gemm_pack_weight(w["down"]), qkv=native.gemm_pack_weight(w["qkv"]))
    ws = native.gemm_chain_workspace(m, [(h, attn.shape[1]), (2 * inter, h), (h, inter), (nq, h)], "cuda")
    r = resid.clone() if in_place else torch.empty_like(resid)
    res_in = r if in_place else resid
    act = torch.empty(m, inter, dtype=BF16, device="cuda")
    qkv = torch.empty(m, nq, dtype=BF16, device="cuda")
    ssq1 = torch.empty(m, h // 128, dtype=torch.float32, device="cuda")
    ssq2 = torch.empty(m, h // 128, dtype=torch.float32, device="cuda")
    ops = [dict(x=attn, packed=packed["o"], n=h, residual=res_in, out=r, epilogue=native.GEMM_RESIDUAL, ssq_out=ssq1),
           dict(x=r, packed=packed["gu"], n=2 * inter, out=act, epilogue=native.GEMM_SILU_MUL, norm_weight=w["ln1"], ssq_in=ssq1),
           dict(x=act, packed=packed["down"], n=h, residual=r, out=r, epilogue=native.GEMM_RESIDUAL, ssq_out=ssq2),
           dict(x=r, packed=packed["qkv"], n=nq, out=qkv, norm_weight=w["ln2"], ssq_in=ssq2)][:nops]
    native.gemm_chain(ops, m, ws, eps=EPS, rounding=rounding)
    return act, r, qkv, ws


SIZES = [(2048, 6144, 4096), (1024, 3072, 4096)]     # (hidden, intermediate, qkv out): 1.7B talker, code predictor


@pytest.mark.parametrize("h,inter,nq", SIZES)
@pytest.mark.parametrize("m", [1, 8, 64, 100])
def test_chain_equals_separate_launches(native, m, h, inter, nq):
    w = _layer_weights(h, inter, nq, seed=m + h)
    g = torch.Generator(device="cuda").manual_seed(5 + m)
    attn = (torch.randn(m, nq // 2, device="cuda", generator=g) * 0.5).to(BF16)
    resid = torch.randn(m, h, device="cuda", generator=g).to(BF16)
    for rounding in (native.ROUND_EAGER, native.ROUND_FUSED):
        r1, act, r2, qkv = _separate(native, w, attn, resid, rounding)
        act_c, r2_c, qkv_c, _ = _chain(native, w, attn, resid, rounding)
        # errors compound along the chain (each op consumes the previous one's rounded output): the activations of
        # the later ops may differ by an ulp wherever an earlier element did
        for name, a, b in (("act", act_c, act), ("residual", r2_c, r2), ("qkv", qkv_c, qkv)):
            rel = ((a.float() - b.float()).abs().max() / b.float().abs().max()).item()
            assert rel < 1e-2, f"{name} m={m}: rel {rel:.3e}"
            frac = (a != b).float().mean().item()          # one-ulp differences compound along the chain
            assert frac < 0.5, f"{name} m={m}: {frac:.3f} of elements differ"


@pytest.mark.parametrize("m", [3, 64])
def test_chain_single_op_and_in_place(native, m):
    """nops = 1 (o_proj + residual only), writing the residual stream in place."""
    h, inter, nq = 2048, 6144, 4096
    w = _layer_weights(h, inter, nq, seed=11)
    g = torch.Generator(device="cuda").manual_seed(m)
    attn = (torch.randn(m, nq // 2, device="cuda", generator=g) * 0.5).to(BF16)
    resid = torch.randn(m, h, device="cuda", generator=g).to(BF16)
    want = (resid.double() + (attn.double() @ w["o"].double().t()).to(BF16).double())
    _, got, _, _ = _chain(native, w, attn, resid, native.ROUND_EAGER, nops=1, in_place=True)
    rel = ((got.double() - want).abs().max() / want.abs().max()).item()
    assert rel < 1e-2, rel


def test_chain_is_deterministic_and_graph_replayable(native):
    """Same bytes on every launch; the barrier counters reset themselves, so a CUDA graph can replay it."""
    m, (h, inter, nq) = 64, SIZES[1]
    w = _layer_weights(h, inter, nq, seed=3)
    g = torch.Generator(device="cuda").manual_seed(9)
    attn = (torch.randn(m, nq // 2, device="cuda", generator=g) * 0.5).to(BF16)
    resid = torch.randn(m, h, device="cuda", generator=g).to(BF16)
    act0, r0, q0, _ = _chain(native, w, attn, resid, native.ROUND_EAGER)
    for _ in range(3):
        act1, r1, q1, ws = _chain(native, w, attn, resid, native.ROUND_EAGER)
        assert torch.equal(act0, act1) and torch.equal(r0, r1) and torch.equal(q0, q1)
    torch.cuda.synchronize()
    assert int(ws[:64].view(torch.int32).abs().sum()) == 0, "barrier counters not reset"
    # graph: 5 chained launches (each consuming the previous residual), replayed 3 times
    packed = dict(o=native.gemm_pack_weight(w["o"]), gu=native.gemm_pack_weight(w["gu"], native.PACK_GATE_UP),
                  down=native.gemm_pack_weight(w["down"]), qkv=native.gemm_pack_weight(w["qkv"]))
    ws = native.gemm_chain_workspace(m, [(h, nq // 2), (2 * inter, h), (h, inter), (nq, h)], "cuda")
    r = resid.clone()
    act = torch.empty(m, inter, dtype=BF16, device="cuda"); qkv = torch.empty(m, nq, dtype=BF16, device="cuda")
    s1 = torch.empty(m, h // 128, dtype=torch.float32, device="cuda"); s2 = torch.empty_like(s1)

    def body():
        for _ in range(5):
            native.gemm_chain([
                dict(x=attn, packed=packed["o"], n=h, residual=r, out=r, epilogue=native.GEMM_RESIDUAL, ssq_out=s1),
                dict(x=r, packed=packed["gu"], n=2 * inter, out=act, epilogue=native.GEMM_SILU_MUL, norm_weight=w["ln1"], ssq_in=s1),
                dict(x=act, packed=packed["down"], n=h, residual=r, out=r, epilogue=native.GEMM_RESIDUAL, ssq_out=s2),
                dict(x=r, packed=packed["qkv"], n=nq, out=qkv, norm_weight=w["ln2"], ssq_in=s2)], m, ws, eps=EPS)
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        body()
    torch.cuda.synchronize()
    want_r, want_q = r.clone(), qkv.clone()
    graph = torch.cuda.CUDAGraph()
    r.copy_(resid)
    with torch.cuda.graph(graph):
        body()
    for _ in range(3):
        r.copy_(resid)
        graph.replay()
        torch.cuda.synchronize()
        assert torch.equal(r, want_r) and torch.equal(qkv, want_q)


def test_chain_argument_errors(native):
    m, (h, inter, nq) = 4, SIZES[1]
    w = _layer_weights(h, inter, nq, seed=1)
    x = torch.zeros(m, nq // 2, dtype=BF16, device="cuda")
    out = torch.zeros(m, h, dtype=BF16, device="cuda")
    ws = native.gemm_chain_workspace(m, [(h, nq // 2)], "cuda")
    with pytest.raises(RuntimeError, match="residual"):
        native.gemm_chain([dict(x=x, packed=native.gemm_pack_weight(w["o"]), n=h, out=out, epilogue=native.GEMM_RESIDUAL)], m, ws)
    with pytest.raises(RuntimeError, match="workspace"):
        native.gemm_chain([dict(x=x, packed=native.gemm_pack_weight(w["o"]), n=h, out=out)], m, ws[:64])
    with pytest.raises(RuntimeError, match="ops"):
        native.gemm_chain([], m, ws)
