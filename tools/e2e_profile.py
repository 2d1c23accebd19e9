"""Host-side profile of the step-API frame (StepApiFrameLoop): cProfile over N frames at B streams.
    python tools/e2e_profile.py [streams] [frames]"""
import cProfile, os, pstats, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from nano_qwen3tts_vllm_b200.engine.frame_loop import StepApiFrameLoop

b = int(sys.argv[1]) if len(sys.argv) > 1 else 64
n = int(sys.argv[2]) if len(sys.argv) > 2 else 24
mc, talker, predictor = bench.build_engines("1.7B", b, n + 16, "cuda:0")
prompts, pad = bench.synthetic_prompts(b, mc.talker.hidden_size, pinned=True)
api = StepApiFrameLoop(talker, predictor, pad)
api.add_streams(prompts)
for _ in range(3):
    api.step_frame()
torch.cuda.synchronize()
api.timing = {}
t0 = time.perf_counter()
for _ in range(n):
    api.step_frame()
torch.cuda.synchronize()
dt = time.perf_counter() - t0
print(f"{b} streams: {dt / n * 1e3:.3f} ms per frame ({b * n / dt:.0f} frames/s); wall ms per frame by phase:")
for k, v in api.timing.items():
    print(f"   {k:45s} {v / n * 1e3:7.3f}")
api.timing = None
t0 = time.perf_counter()
pr = cProfile.Profile()
pr.enable()
for _ in range(n):
    api.step_frame()
torch.cuda.synchronize()
pr.disable()
dt = time.perf_counter() - t0
print(f"{b} streams: {dt / n * 1e3:.3f} ms per frame under cProfile ({b * n / dt:.0f} frames/s)")
pstats.Stats(pr).sort_stats("tottime").print_stats(22)
