"""Kernel micro-benchmarks (CUDA events on the launching stream, cold L2 by rotating among
cache copies larger than the 126 MB L2).  Prints one JSON line per case.

    python tools/kbench.py attn   [--cases B:L,B:L,...]
    python tools/kbench.py kvwrite|norm|silu|sampler|qkrope
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nano_qwen3tts_vllm_b200._native as nat  # noqa: E402

BF16 = torch.bfloat16
DEV = "cuda"


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        return 6650.0


def time_ms(fn, iters=20, warm=5):
    for _ in range(warm):
        fn(0)
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for i, (a, b) in enumerate(evs):
        a.record()
        fn(i)
        b.record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in evs)
    return ts[len(ts) // 2]


def attn_case(b, l, ragged=False, page=256, kvh=8, h=16, hd=128):
    g = torch.Generator().manual_seed(11)
    lens = torch.randint(64, l + 1, (b,), generator=g) if ragged else torch.full((b,), l)
    pages_per = (l + page - 1) // page
    npages = b * pages_per
    bytes_per_copy = npages * page * kvh * hd * 2 * 2
    ncopies = max(2, min(8, int(300e6 // bytes_per_copy) + 1))
    caches = [(torch.randn(npages, page, kvh, hd, device=DEV, dtype=BF16),
               torch.randn(npages, page, kvh, hd, device=DEV, dtype=BF16)) for _ in range(ncopies)]
    perm = torch.randperm(npages, generator=torch.Generator().manual_seed(13)).to(torch.int32)
    bt = perm.view(b, pages_per).to(DEV)
    cl = lens.to(torch.int32).to(DEV)
    q = torch.randn(b, h, hd, device=DEV, dtype=BF16)
    max_ctx = pages_per * page
    ws = torch.empty(max(16, nat.paged_attn_decode_workspace_bytes(b, h, hd, max_ctx)), dtype=torch.uint8, device=DEV)
    out = torch.empty(b, h, hd, device=DEV, dtype=BF16)

    def fn(i):
        kc, vc = caches[i % ncopies]
        nat.paged_attn_decode(q, kc, vc, bt, cl, hd ** -0.5, max_ctx, out=out, workspace=ws)

    ms = time_ms(fn)
    alg = float((4096 * lens + 8192).sum())
    gbs = alg / ms / 1e6
    print(json.dumps({"kernel": "paged_attn_decode", "B": b, "L": l, "ragged": ragged, "ms": round(ms, 5),
                      "alg_bytes": alg, "GBps": round(gbs, 1), "frac_measured_peak": round(gbs / peaks(), 4)}))


def gemm_case(m, n, k, splits=0, pdl=False, epilogue=0, chain=16):
    """`chain` GEMMs back to back over DIFFERENT weights (a stack of layers: cold weights, > L2 in
    total), timed as one CUDA graph so that launch gaps are what a decode step sees."""
    ncopies = max(chain, int(300e6 // (n * k * 2)) + 1)
    ws = [(torch.randn(n, k, device=DEV, dtype=torch.float32) * 0.05).to(BF16) for _ in range(ncopies)]
    packed = [nat.gemm_pack_weight(w, nat.PACK_GATE_UP if epilogue == 2 else nat.PACK_PLAIN) for w in ws]
    x = torch.randn(m, k, device=DEV, dtype=BF16)
    res = torch.randn(m, n, device=DEV, dtype=BF16) if epilogue == 1 else None
    out = torch.empty(m, n // 2 if epilogue == 2 else n, device=DEV, dtype=BF16)
    rows = []
    for impl in ("ours", "cublas"):
        def body(i0):
            for j in range(chain):
                i = (i0 * chain + j) % ncopies
                if impl == "ours":
                    nat.gemm_bf16(x, packed[i], n, residual=res, epilogue=epilogue, out=out, splits=splits, pdl=pdl)
                else:
                    torch.nn.functional.linear(x, ws[i])
        body(0)
        torch.cuda.synchronize()
        graphs = []
        for i0 in range(max(1, ncopies // chain)):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                body(i0)
            graphs.append(g)
        ms = time_ms(lambda i: graphs[i % len(graphs)].replay()) / chain
        alg = float(n * k * 2 + m * k * 2 + m * n * 2)
        rows.append((impl, ms, alg / ms / 1e6))
    (_, ms_o, gb_o), (_, ms_c, gb_c) = rows
    print(json.dumps({"kernel": "gemm_bf16", "M": m, "N": n, "K": k, "splits": splits, "pdl": pdl, "epilogue": epilogue,
                      "us": round(ms_o * 1e3, 2), "GBps": round(gb_o, 1), "frac_measured_peak": round(gb_o / peaks(), 4),
                      "cublas_us": round(ms_c * 1e3, 2), "cublas_GBps": round(gb_c, 1)}), flush=True)


def sampler_case(b, v, top_k=50, reps=16):
    """`reps` top-k sampler launches per graph over different logits rows (lm_head output shape [b, v] bf16)."""
    logits = [(torch.randn(b, v, device=DEV) * 2).to(BF16) for _ in range(reps)]
    noise = torch.empty(b, v, device=DEV, dtype=torch.float32).exponential_(1)
    temps = torch.full((b,), 0.9, device=DEV, dtype=torch.float32)
    out = torch.empty(b, device=DEV, dtype=torch.int64)

    def fn():
        for lg in logits:
            nat.topk_sample(lg, temps, noise, top_k=top_k, out=out)
    fn(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    us = time_ms(lambda i: g.replay()) / reps * 1e3
    alg = float(b * (2 * v + 4 * v + 12))
    print(json.dumps({"kernel": "topk_sample", "B": b, "V": v, "top_k": top_k, "us": round(us, 2),
                      "GBps": round(alg / us / 1e3, 1)}), flush=True)


def chain_case(m, h, inter, nq, layers=8):
    """`layers` GEMM chains (o + gate_up + down + next qkv) back to back over different weights in one graph,
    against the same 4 x layers projections as separate b2tts_gemm_norm_bf16 launches."""
    g = torch.Generator(device=DEV).manual_seed(1)
    mk = lambda n, k: (torch.randn(n, k, device=DEV, generator=g) * 0.03).to(BF16)
    L = []
    for _ in range(layers):
        L.append(dict(o=nat.gemm_pack_weight(mk(h, nq // 2)), gu=nat.gemm_pack_weight(mk(2 * inter, h), nat.PACK_GATE_UP),
                      down=nat.gemm_pack_weight(mk(h, inter)), qkv=nat.gemm_pack_weight(mk(nq, h)),
                      ln1=torch.ones(h, device=DEV, dtype=BF16), ln2=torch.ones(h, device=DEV, dtype=BF16)))
    attn = torch.randn(m, nq // 2, device=DEV, dtype=BF16)
    r = torch.randn(m, h, device=DEV, dtype=BF16)
    act = torch.empty(m, inter, device=DEV, dtype=BF16); qkv = torch.empty(m, nq, device=DEV, dtype=BF16)
    s1 = torch.empty(m, h // 128, device=DEV, dtype=torch.float32); s2 = torch.empty_like(s1)
    ws = nat.gemm_chain_workspace(m, [(h, nq // 2), (2 * inter, h), (h, inter), (nq, h)], DEV)
    t1 = torch.empty(m, nat.gemm_ssq_slots(m, h, nq // 2), device=DEV, dtype=torch.float32)
    t2 = torch.empty(m, nat.gemm_ssq_slots(m, h, inter), device=DEV, dtype=torch.float32)

    def chained():
        for w in L:
            nat.gemm_chain([
                dict(x=attn, packed=w["o"], n=h, residual=r, out=r, epilogue=nat.GEMM_RESIDUAL, ssq_out=s1),
                dict(x=r, packed=w["gu"], n=2 * inter, out=act, epilogue=nat.GEMM_SILU_MUL, norm_weight=w["ln1"], ssq_in=s1),
                dict(x=act, packed=w["down"], n=h, residual=r, out=r, epilogue=nat.GEMM_RESIDUAL, ssq_out=s2),
                dict(x=r, packed=w["qkv"], n=nq, out=qkv, norm_weight=w["ln2"], ssq_in=s2)], m, ws)

    def separate():
        for w in L:
            nat.gemm_bf16(attn, w["o"], h, residual=r, epilogue=nat.GEMM_RESIDUAL, out=r, ssq_out=t1, pdl=True)
            nat.gemm_bf16(r, w["gu"], 2 * inter, epilogue=nat.GEMM_SILU_MUL, out=act, norm_weight=w["ln1"], ssq_in=t1, pdl=True)
            nat.gemm_bf16(act, w["down"], h, residual=r, epilogue=nat.GEMM_RESIDUAL, out=r, ssq_out=t2, pdl=True)
            nat.gemm_bf16(r, w["qkv"], nq, out=qkv, norm_weight=w["ln2"], ssq_in=t2, pdl=True)
    res = {}
    for name, fn in (("chain", chained), ("separate", separate)):
        fn(); torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            fn()
        res[name] = time_ms(lambda i: gr.replay()) / layers * 1e3
    alg = float(2 * (h * nq // 2 + 2 * inter * h + h * inter + nq * h))
    print(json.dumps({"kernel": "gemm_chain", "M": m, "hidden": h, "inter": inter, "qkv": nq, "chain_us": round(res["chain"], 2),
                      "separate_us": round(res["separate"], 2), "chain_GBps": round(alg / res["chain"] / 1e3, 1),
                      "frac_measured_peak": round(alg / res["chain"] / 1e3 / peaks(), 4),
                      "env": {k: v for k, v in os.environ.items() if k.startswith("B2TTS_")}}), flush=True)


def fattn_case(b, l, layers=28, page=256, kvh=8, h=16, hd=128):
    """Fused decode attention vs the three separate launches, `layers` launches per graph over different
    caches (cold KV), uniform context l."""
    pages_per = (l + page - 1) // page
    npages = b * pages_per
    caches = [(torch.randn(npages, page, kvh, hd, device=DEV, dtype=BF16),
               torch.randn(npages, page, kvh, hd, device=DEV, dtype=BF16)) for _ in range(layers)]
    bt = torch.randperm(npages, generator=torch.Generator().manual_seed(13)).to(torch.int32).view(b, pages_per).to(DEV)
    lens = torch.full((b,), l, dtype=torch.int32, device=DEV)
    slots = (bt[:, (l - 1) // page] * page + (l - 1) % page).to(torch.int32).contiguous()
    qkv = torch.randn(b, (h + 2 * kvh) * hd, device=DEV, dtype=BF16)
    q, k, v = qkv[:, :h * hd].view(b, h, hd), qkv[:, h * hd:(h + kvh) * hd].view(b, kvh, hd), qkv[:, (h + kvh) * hd:].view(b, kvh, hd)
    w = torch.ones(hd, device=DEV, dtype=BF16)
    inv_freq = (1.0 / (1e6 ** (torch.arange(0, hd, 2, device=DEV).float() / hd))).contiguous()
    pos = (lens.long() - 1).contiguous()
    max_ctx = pages_per * page
    ws = nat.fused_decode_attn_workspace(b, h, hd, max_ctx, DEV)
    ws2 = torch.empty(max(16, nat.paged_attn_decode_workspace_bytes(b, h, hd, max_ctx)), dtype=torch.uint8, device=DEV)
    out = torch.empty(b, h, hd, device=DEV, dtype=BF16)

    def fused():
        for kc, vc in caches:
            nat.fused_decode_attn(q, k, v, w, w, 1e-6, pos, inv_freq, slots, kc, vc, bt, lens, hd ** -0.5, max_ctx, ws, out=out)

    def separate():
        for kc, vc in caches:
            qr, _ = nat.qknorm_rope_kvwrite(q, k, v, w, w, 1e-6, pos, inv_freq, slot_mapping=slots, k_cache=kc, v_cache=vc,
                                            want_k_out=False)
            nat.paged_attn_decode(qr, kc, vc, bt, lens, hd ** -0.5, max_ctx, out=out, workspace=ws2)
    res = {}
    for name, fn in (("fused", fused), ("separate", separate)):
        fn(); torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fn()
        res[name] = time_ms(lambda i: g.replay()) / layers * 1e3
    alg = float(b * (4096 * l + 8192))
    print(json.dumps({"kernel": "fused_decode_attn", "B": b, "L": l, "fused_us": round(res["fused"], 2),
                      "separate_us": round(res["separate"], 2), "fused_GBps": round(alg / res["fused"] / 1e3, 1),
                      "nsplit_env": os.environ.get("B2TTS_FA_NSPLIT")}), flush=True)


def prefill_case(lens, h=16, kvh=8, hd=128, layers=4):
    """Varlen causal prefill attention (a3) over packed q/k/v slices of one QKV projection: ours (tcgen05 tile kernel
    for max_seqlen >= 128, B2TTS_PREFILL_TC=0 selects the per-query kernel) and flash_attn_varlen_func."""
    t = sum(lens)
    qkvs = [torch.randn(t, (h + 2 * kvh) * hd, device=DEV, dtype=BF16) for _ in range(layers)]
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32, device=DEV)
    out = torch.empty(t, h, hd, device=DEV, dtype=BF16)
    scale = hd ** -0.5

    def split(x):
        return x[:, :h * hd].view(t, h, hd), x[:, h * hd:(h + kvh) * hd].view(t, kvh, hd), x[:, (h + kvh) * hd:].view(t, kvh, hd)

    def ours():
        for x in qkvs:
            q, k, v = split(x)
            nat.varlen_attn_prefill(q, k, v, cu, cu, max(lens), max(lens), scale, out=out)
    res = {}
    impls = [("ours", ours)]
    try:
        import flash_attn

        def fa():
            for x in qkvs:
                q, k, v = split(x)
                flash_attn.flash_attn_varlen_func(q, k, v, cu_seqlens_q=cu, cu_seqlens_k=cu, max_seqlen_q=max(lens),
                                                  max_seqlen_k=max(lens), softmax_scale=scale, causal=True)
        impls.append(("flash_attn", fa))
    except ImportError:
        pass
    for name, fn in impls:
        fn(); torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fn()
        res[name] = time_ms(lambda i: g.replay()) / layers * 1e3
    flops = sum(2.0 * n * n * hd * h for n in lens)          # causal: half of 4 * T^2 * hd * heads
    tf = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("bf16_tflops_sustained", 1378.0) \
        if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 1378.0
    row = {"kernel": "varlen_attn_prefill", "lens": lens if len(lens) <= 4 else f"{len(lens)}x{lens[0]}", "T": t,
           "tc": os.environ.get("B2TTS_PREFILL_TC", "1"), "us": round(res["ours"], 2),
           "TFLOPs": round(flops / res["ours"] / 1e6, 1), "frac_sustained_bf16": round(flops / res["ours"] / 1e6 / tf, 4)}
    if "flash_attn" in res:
        row.update(flash_attn_us=round(res["flash_attn"], 2), flash_attn_TFLOPs=round(flops / res["flash_attn"] / 1e6, 1))
    print(json.dumps(row), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("what")
    ap.add_argument("--cases", default="1:512,8:512,64:512,64:1024,256:512,512:600,8:4352,64:4096")
    ap.add_argument("--ragged", action="store_true")
    ap.add_argument("--gemm-cases", default="64:4096:2048,64:2048:2048,64:12288:2048,64:2048:6144,64:4096:1024,"
                                            "64:1024:2048,64:6144:1024,64:1024:3072,64:2048:1024,1:4096:2048,8:4096:2048,"
                                            "256:4096:2048")
    ap.add_argument("--chain-cases", default="64:2048:6144:4096,64:1024:3072:4096,8:2048:6144:4096,1:2048:6144:4096")
    ap.add_argument("--splits", default="0")
    ap.add_argument("--pdl", type=int, default=1)
    a = ap.parse_args()
    if a.what == "gemm":
        for c in a.gemm_cases.split(","):
            m, n, k = map(int, c.split(":"))
            for s_ in [int(v) for v in a.splits.split(",")]:
                if s_ and (s_ > k // 64):
                    continue
                gemm_case(m, n, k, splits=s_, pdl=bool(a.pdl))
    if a.what == "prefill":
        for lens in ([4096], [1024], [256], [4096, 4096], [64] * 64, [300, 5]):
            prefill_case(lens)
    if a.what == "sampler":
        for b, v in ((64, 2048), (64, 3072), (1, 3072), (512, 2048)):
            sampler_case(b, v)
    if a.what == "chain":
        for c in a.chain_cases.split(","):
            m, h, inter, nq = map(int, c.split(":"))
            chain_case(m, h, inter, nq)
    if a.what == "fattn":
        for c in a.cases.split(","):
            b, l = map(int, c.split(":"))
            fattn_case(b, l, layers=5 if l <= 32 else 28)
    if a.what == "attn":
        for c in a.cases.split(","):
            b, l = map(int, c.split(":"))
            attn_case(b, l, a.ragged)


if __name__ == "__main__":
    main()
