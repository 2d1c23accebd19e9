"""cProfile of the end-to-end arm (StepApiFrameLoop through TalkerLLM / PredictorLLM with host buffers): where the host
time of one frame goes.   python tools/profile_e2e.py [streams] [frames]"""
import cProfile, os, pstats, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench

b = int(sys.argv[1]) if len(sys.argv) > 1 else 64
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 16
mc, talker, predictor = bench.build_engines("1.7B", b, frames + 16, "cuda:0")
bench.run_e2e(talker, predictor, mc, b, 4)                 # warm
pr = cProfile.Profile()
pr.enable()
fps, h2d, d2h, dt = bench.run_e2e(talker, predictor, mc, b, frames)
pr.disable()
print(f"e2e {fps:.1f} frames/s, {1e3 * dt / frames:.2f} ms per frame of {b} streams")
pstats.Stats(pr).sort_stats("cumulative").print_stats(45)
