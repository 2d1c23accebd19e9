"""In-frame timeline of the talker's GEMM launches (debug): two decoder layers of the real model in one CUDA graph
(QKV GEMM, fused attention, o_proj, gate_up, down per layer), every GEMM CTA stamping %globaltimer; the gaps between
one GEMM's end and the next one's dependency resolution are the attention kernel / launch boundaries.
    python tools/layer_trace.py [streams] [context]"""
import ctypes, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import nano_qwen3tts_vllm_b200._native as nat
from nano_qwen3tts_vllm_b200.engine.frame_loop import DeviceFrameLoop

b = int(sys.argv[1]) if len(sys.argv) > 1 else 64
dev = "cuda:0"
mc, talker, predictor = bench.build_engines("1.7B", b, 64, dev)
loop = DeviceFrameLoop(talker.model_runner, predictor.model_runner, b, 40, bench.PROMPT_LEN)
prompts, pad = bench.synthetic_prompts(b, mc.talker.hidden_size, pinned=False)
loop.start([p[0].to(dev) for p in prompts], pad.to(dev))
for _ in range(3):
    loop.step()
torch.cuda.synchronize()
from nano_qwen3tts_vllm_b200.utils.context import reset_context, set_context
WHICH = sys.argv[2] if len(sys.argv) > 2 else "talker"          # "talker" | "predictor"
T = loop.talker if WHICH == "talker" else loop.predictor
layers = T.model.layers
NL = 3
lib = nat.lib()
lib.b2tts_debug_gemm_trace.argtypes = [ctypes.c_void_p, ctypes.c_int]


def body():
    if WHICH == "talker":
        set_context(False, slot_mapping=loop.t_slots, context_lens=loop.t_ctx, block_tables=loop.t_tables,
                    max_context_len=loop.width * loop.page)
        residual, ssq, pos = loop.embeds.clone(), None, loop.t_pos
    else:                                                       # the predictor's decode pass for codebook 5
        lay = loop.p_dec[4]
        set_context(False, slot_mapping=lay["slots"], context_lens=lay["ctx"], block_tables=loop.p_tables, max_context_len=7)
        residual, ssq, pos = torch.randn(b, layers[0].input_layernorm.weight.shape[0], device=dev).to(torch.bfloat16), None, lay["pos"]
    for i in range(NL):
        residual, ssq = layers[i].forward_decode(pos, residual, ssq, in_place=True)
    reset_context()


with torch.inference_mode():
    nat.decode_layout(loop.seq_lens, loop.t_tables, loop.page, 1, 0, loop.t_pos, loop.t_slots, loop.t_ctx)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        body()
    torch.cuda.synchronize()
    nl = 4 * NL
    trace = torch.zeros(nl * 4096 * 8, dtype=torch.int64, device=dev)
    lib.b2tts_debug_gemm_trace(trace.data_ptr(), nl)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        body()
    lib.b2tts_debug_gemm_trace(None, 0)
with torch.inference_mode():
    g.replay(); torch.cuda.synchronize()
    trace.zero_()
    g.replay(); torch.cuda.synchronize()
t = trace.view(nl, 4096, 8).cpu()
names = ["entry", "setup", "dep ok", "x first", "accum", "staged", "exchanged", "end"]
ops = ["qkv(+rmsnorm before layer 0)", "o_proj", "gate_up", "down"]
base = None
print(f"streams={b}; us since the first stamp; min/median/max over the launch's CTAs; first launch of layer 0 is preceded by an rmsnorm")
prev_end = None
for l in range(nl):
    n = int((t[l, :, 0] != 0).sum())
    if n == 0:
        print(f"launch {l}: no stamps"); continue
    if base is None:
        base = int(t[l, :n, 0].min())
    tt = (t[l, :n].double() - base) / 1e3
    gap = "" if prev_end is None else f"  [gap from previous GEMM end to dep ok: {float(tt[:, 2].min()) - prev_end:5.1f}]"
    print(f"{ops[l % 4]:30s} CTAs {n:3d} " + "  ".join(f"{names[j]} {tt[:, j].min():6.1f}/{tt[:, j].median():6.1f}/{tt[:, j].max():6.1f}" for j in (0, 2, 4, 6, 7)) + gap)
    per = tt[:, 5] - tt[:, 4]; ex = tt[:, 6] - tt[:, 5]; ep = tt[:, 7] - tt[:, 6]; xf = tt[:, 3] - tt[:, 2]; mm = tt[:, 4] - tt[:, 3]
    print(f"{'':30s} per CTA (median): dep ok->first x tile released {xf.median():.2f}  ->accum {mm.median():.2f}  accum->staged {per.median():.2f}  staged->exchanged {ex.median():.2f} (min {ex.min():.2f} max {ex.max():.2f})  exchanged->end {ep.median():.2f}")
    prev_end = float(tt[:, 7].max())
