"""Fused decode attention at the contexts a frame starts with: launch time against the split-KV factor
(B2TTS_FA_NSPLIT, experiments only).   python tools/attn_short_ctx.py"""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 1 and sys.argv[1] == "--child":
    sys.path.insert(0, ROOT)
    import torch
    import bench_kernels as BK
    for b in (64, 256):
        for ctx in (65, 130, 260, 520):
            ms, alg = BK.decode_attention(b, ctx, "cuda:0", layers=8)
            print(f"  streams {b:3d} ctx {ctx:4d}: {ms * 1e3:7.2f} us  {alg / ms / 1e6:7.1f} GB/s", flush=True)
    sys.exit(0)
for ns in ("", "1", "2", "3", "4", "6"):
    env = dict(os.environ)
    if ns:
        env["B2TTS_FA_NSPLIT"] = ns
    print(f"nsplit {ns or 'auto'}", flush=True)
    subprocess.run([sys.executable, __file__, "--child"], env=env)
