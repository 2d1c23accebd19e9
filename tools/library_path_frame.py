"""Frame time of the device frame loop with the projections on the hand-written kernel (fused norm / residual / SiLU*mul
epilogues, 5 launches per layer) against the same loop with every projection on the library GEMM (F.linear = cuBLAS) and
the norms / residual adds / SiLU*mul as separate launches of this repo's kernels (B2TTS_GEMM_NATIVE_MAX_M=0) - the
layer-level answer to "the library GEMM is faster per launch".   python tools/library_path_frame.py [streams] [frames]"""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 1 and sys.argv[1] == "--child":
    sys.path.insert(0, ROOT)
    import bench
    streams, frames = int(sys.argv[2]), int(sys.argv[3])
    mc, talker, predictor = bench.build_engines("1.7B", streams, frames + 32, "cuda:0")
    loop, ms, launches, _ = bench.run_device_loop(talker, predictor, mc, streams, frames, 8, None)
    print(json.dumps({"ms_per_frame": ms / frames, "launches_per_frame": launches}))
    sys.exit(0)
streams = int(sys.argv[1]) if len(sys.argv) > 1 else 64
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 96
for label, env in (("hand-written GEMM, fused epilogues", {}), ("library GEMM (cuBLAS), separate norm / add / SiLU*mul launches", {"B2TTS_GEMM_NATIVE_MAX_M": "0"})):
    out = subprocess.run([sys.executable, __file__, "--child", str(streams), str(frames)], env={**os.environ, **env},
                         capture_output=True, text=True)
    line = [l for l in out.stdout.splitlines() if l.startswith("{")]
    print(f"{streams} streams, {label}: {line[-1] if line else out.stderr[-400:]}")
