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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("what")
    ap.add_argument("--cases", default="1:512,8:512,64:512,64:1024,256:512,512:600,8:4352,64:4096")
    ap.add_argument("--ragged", action="store_true")
    a = ap.parse_args()
    if a.what == "attn":
        for c in a.cases.split(","):
            b, l = map(int, c.split(":"))
            attn_case(b, l, a.ragged)


if __name__ == "__main__":
    main()
