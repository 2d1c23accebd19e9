"""Greedy (top_k=1) frame-loop identity: DeviceFrameLoop / StepApiFrameLoop vs the CPU oracle's run_frames.
Prints the first diverging (frame, stream, codebook) and the oracle's logit margin there."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import model as OM, ops as O
from tests.test_gpu_models import small_model_config, small_weights
from nano_qwen3tts_vllm_b200.llm import PredictorLLM, TalkerLLM
from nano_qwen3tts_vllm_b200.engine.frame_loop import DeviceFrameLoop, StepApiFrameLoop

BF16 = torch.bfloat16
nframes = int(sys.argv[1]) if len(sys.argv) > 1 else 8
b = int(sys.argv[2]) if len(sys.argv) > 2 else 2
td, pd, tw, pw = small_weights()
g = torch.Generator().manual_seed(1000)
prompts = [(torch.randn(n, td.hidden_size, generator=g) * 0.05).to(BF16) for n in ([9, 14, 5, 11] * 4)[:b]]
pad = (torch.randn(1, td.hidden_size, generator=torch.Generator().manual_seed(999)) * 0.05).to(BF16)

margins = []
def greedy(logits, temps):
    top2 = logits.float().topk(2, dim=-1).values
    margins.append((top2[:, 0] - top2[:, 1], top2[:, 0]))
    return O.topk_sample(logits, temps, torch.ones(logits.shape), top_k=1)
def margin_at(div):
    f, s_, c = div
    m, top = margins[16 * f + c]
    return f"oracle top-2 margin {float(m[s_]):.4g} at logit {float(top[s_]):.4g} (bf16 ulp there {2.0 ** (torch.tensor(float(top[s_])).abs().log2().floor().item() - 7):.4g})"
ot = OM.Talker(td, tw, b * 2)
op = OM.Predictor(pd, pw, b, td.hidden_size)
want = OM.run_frames(ot, op, prompts, nframes, pad, greedy)
want_t = torch.tensor(want.frames)                     # [frames, b, 16]

mc = small_model_config()
kw = dict(num_kvcache_blocks=2 * b, max_num_seqs=max(b, 2), max_model_len=512, enforce_eager=True)
talker = TalkerLLM(mc, state_dict=tw, **kw)
predictor = PredictorLLM(mc, state_dict=pw, **kw)
loop = DeviceFrameLoop(talker.model_runner, predictor.model_runner, b, nframes, 16, top_k=1, scatter_pages=False)
loop.start([p.cuda() for p in prompts], pad.cuda())
loop.capture()
loop.run(nframes)
got = loop.result()
def first_div(a, w):
    ne = (a != w).nonzero()
    return None if ne.numel() == 0 else tuple(ne[0].tolist())
d = first_div(got, want_t)
print("device loop vs oracle: first divergence", d, "of", tuple(want_t.shape), margin_at(d) if d else "")
api = StepApiFrameLoop(talker, predictor, pad)
talker.model_runner.top_k = predictor.model_runner.top_k = 1
talker.model_runner.kv_cache.zero_(); predictor.model_runner.kv_cache.zero_()
api.add_streams([p.unsqueeze(0).pin_memory() for p in prompts])
for _ in range(nframes):
    api.step_frame()
got2 = torch.tensor([api.frames[f"s{i}"] for i in range(b)]).permute(1, 0, 2)
d = first_div(got2, want_t)
print("step-API loop vs oracle: first divergence", d, margin_at(d) if d else "")
print("step-API loop vs device loop: first divergence", first_div(got2, got))
print("h2d/d2h bytes per frame:", api.h2d_bytes / nframes, api.d2h_bytes / nframes)
