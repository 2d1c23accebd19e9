"""What each kernel family costs INSIDE a captured pass (not in isolation): a talker / predictor decode pass is
captured as a CUDA graph, then captured again with one family of launches replaced by a no-op that returns a
pre-allocated tensor; the difference of the two replay times is that family's share of the dependent chain.
    python tools/pass_ablation.py [streams]"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import nano_qwen3tts_vllm_b200._native as nat
from nano_qwen3tts_vllm_b200.engine.frame_loop import DeviceFrameLoop
from nano_qwen3tts_vllm_b200.layers import linear as L
from nano_qwen3tts_vllm_b200.models import qwen3_tts_share as S
from nano_qwen3tts_vllm_b200.utils.context import reset_context, set_context

b = int(sys.argv[1]) if len(sys.argv) > 1 else 64
dev = "cuda:0"
mc, talker, predictor = bench.build_engines("1.7B", b, 64, dev)
loop = DeviceFrameLoop(talker.model_runner, predictor.model_runner, b, 40, bench.PROMPT_LEN)
prompts, pad = bench.synthetic_prompts(b, mc.talker.hidden_size, pinned=False)
loop.start([p[0].to(dev) for p in prompts], pad.to(dev))
for _ in range(3):
    loop.step()
torch.cuda.synchronize()
T, P = loop.talker, loop.predictor


def timed(fn, reps=30):
    with torch.inference_mode():
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            fn()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fn()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    return 1e3 * e0.elapsed_time(e1) / reps


def talker_pass():
    nat.decode_layout(loop.seq_lens, loop.t_tables, loop.page, 1, 0, loop.t_pos, loop.t_slots, loop.t_ctx)
    set_context(False, slot_mapping=loop.t_slots, context_lens=loop.t_ctx, block_tables=loop.t_tables,
                max_context_len=loop.width * loop.page)
    h = T(loop.embeds, loop.t_pos); reset_context(); return h


def pred_pass(g=5):
    lay = loop.p_dec[g - 1]
    set_context(False, slot_mapping=lay["slots"], context_lens=lay["ctx"], block_tables=loop.p_tables, max_context_len=g + 2)
    h = P(loop.p_emb, lay["pos"])
    loop._sample(L.linear(h, P.lm_head[g]), loop.p_temp, g + 1)
    reset_context(); return h


class Patch:
    def __init__(self, obj, name, fn):
        self.obj, self.name, self.fn = obj, name, fn
    def __enter__(self):
        self.old = getattr(self.obj, self.name); setattr(self.obj, self.name, self.fn)
    def __exit__(self, *a):
        setattr(self.obj, self.name, self.old)


_cache = {}
def cached(key, make):
    if key not in _cache:
        _cache[key] = make()
    return _cache[key]


def no_attn(q, k, v, *a, **kw):
    return cached(("attn", q.shape), lambda: torch.zeros(q.shape, dtype=q.dtype, device=q.device))


real_gemm = nat.gemm_bf16
def no_gemm(x, packed, n, **kw):
    out = kw.get("out")
    cols = n // 2 if kw.get("epilogue") == nat.GEMM_SILU_MUL else n
    if out is None:
        out = cached(("gemm", x.shape[0], cols), lambda: torch.zeros(x.shape[0], cols, dtype=x.dtype, device=x.device))
    so = kw.get("ssq_out")
    return out


def no_sample(logits, temps, noise, **kw):
    return kw.get("out")


for name, fn in (("talker decode pass", talker_pass), ("predictor decode pass + head + sampler", pred_pass)):
    full = timed(fn)
    with Patch(nat, "fused_decode_attn", no_attn):
        wo_attn = timed(fn)
    with Patch(nat, "gemm_bf16", no_gemm):
        wo_gemm = timed(fn)
    with Patch(nat, "topk_sample", no_sample):
        wo_samp = timed(fn)
    with Patch(nat, "gemm_bf16", no_gemm), Patch(nat, "fused_decode_attn", no_attn), Patch(nat, "topk_sample", no_sample):
        rest = timed(fn)
    print(f"{name:42s} full {full:8.1f} us | attention {full - wo_attn:7.1f} | GEMMs {full - wo_gemm:7.1f} | "
          f"sampler {full - wo_samp:6.1f} | everything else {rest:6.1f}", flush=True)
