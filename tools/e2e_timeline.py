"""Where the step-API frame's wall time goes on the host: wall time per wrapped function (the .tolist() calls are the
waits for the GPU; everything else is host work during which the GPU idles).
    python tools/e2e_timeline.py [streams] [frames]"""
import os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from nano_qwen3tts_vllm_b200.engine import frame_loop as fl

b = int(sys.argv[1]) if len(sys.argv) > 1 else 64
n = int(sys.argv[2]) if len(sys.argv) > 2 else 48
mc, talker, predictor = bench.build_engines("1.7B", b, n + 16, "cuda:0")
prompts, pad = bench.synthetic_prompts(b, mc.talker.hidden_size, pinned=True)
api = fl.StepApiFrameLoop(talker, predictor, pad)
api.add_streams(prompts)
for _ in range(4):
    api.step_frame()
torch.cuda.synchronize()

acc, depth = {}, [0]


def wrap(obj, name, label):
    fn = getattr(obj, name)

    def inner(*a, **k):
        t0 = time.perf_counter()
        try:
            return fn(*a, **k)
        finally:
            acc[label] = acc.get(label, 0.0) + time.perf_counter() - t0
    setattr(obj, name, inner)


tr, pr = talker.model_runner, predictor.model_runner
wrap(talker.scheduler, "schedule", "talker  scheduler.schedule")
wrap(tr, "fill_decode_layout", "talker  fill_decode_layout (+upload)")
wrap(tr, "run_model", "talker  run_model (layout, embeds copy, graph launch, head)")
wrap(tr, "run", "talker  run (run_model + sampler + tolist + clone)")
wrap(talker.scheduler, "postprocess", "talker  scheduler.postprocess")
wrap(talker, "add_requests", "talker  add_requests")
wrap(predictor, "add_requests", "predictor add_requests")
wrap(predictor.scheduler, "schedule", "predictor scheduler.schedule")
wrap(pr, "run_burst", "predictor run_burst (layout, upload, x copy, graph launch)")
wrap(predictor, "_step_burst", "predictor _step_burst (run_burst + finish + tolist)")
wrap(api, "_to_host", "loop    _to_host (D2H enqueue)")
wrap(fl.nat, "gather_rows", "loop    gather_rows launch")
wrap(fl.nat, "codec_embed_sum", "loop    codec_embed_sum launch")
orig_sync = torch.cuda.Stream.synchronize


def sync(self):
    t0 = time.perf_counter()
    orig_sync(self)
    acc["loop    stream.synchronize"] = acc.get("loop    stream.synchronize", 0.0) + time.perf_counter() - t0


torch.cuda.Stream.synchronize = sync
api.timing = {}
t0 = time.perf_counter()
for _ in range(n):
    api.step_frame()
torch.cuda.synchronize()
dt = time.perf_counter() - t0
print(f"{b} streams: {dt / n * 1e3:.3f} ms per frame ({b * n / dt:.0f} frames/s)")
for k, v in api.timing.items():
    print(f"   phase {k:50s} {v / n * 1e3:7.3f} ms")
for k, v in sorted(acc.items()):
    print(f"   {k:62s} {v / n * 1e3:7.3f} ms")
