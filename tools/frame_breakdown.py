"""Where one codec frame goes: the pieces of DeviceFrameLoop._frame captured as separate CUDA graphs and
timed with CUDA events (each graph replayed alone, warm).  python tools/frame_breakdown.py [streams]"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import nano_qwen3tts_vllm_b200._native as nat
from nano_qwen3tts_vllm_b200.engine.frame_loop import DeviceFrameLoop
from nano_qwen3tts_vllm_b200.layers.linear import linear
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


def timed(name, fn, reps=20):
    with torch.inference_mode():
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            n0 = nat.launch_count(); fn(); n1 = nat.launch_count()
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
    print(f"{name:34s} {1e3 * e0.elapsed_time(e1) / reps:9.1f} us   ({n1 - n0} lib launches)", flush=True)


def talker_pass():
    nat.decode_layout(loop.seq_lens, loop.t_tables, loop.page, 1, 0, loop.t_pos, loop.t_slots, loop.t_ctx)
    set_context(False, slot_mapping=loop.t_slots, context_lens=loop.t_ctx, block_tables=loop.t_tables,
                max_context_len=loop.width * loop.page)
    h = T(loop.embeds, loop.t_pos); reset_context(); return h

def pred_decode(g=5):
    lay = loop.p_dec[g - 1]
    emb = nat.gather_rows(P.model.codec_embedding[g - 1], loop.codes_t[g])
    set_context(False, slot_mapping=lay["slots"], context_lens=lay["ctx"], block_tables=loop.p_tables, max_context_len=g + 2)
    h = P(emb, lay["pos"]); reset_context(); return h

def pred_prefill():
    set_context(True, loop.p_cu, loop.p_cu, 2, 2, loop.p_pre_slots, None, None)
    h = P(loop.p_x, loop.p_pre_pos); reset_context(); return h

hid_p = pred_decode()
def head_sample():
    loop._sample(linear(hid_p, P.lm_head[5]), loop.p_temp, 6)

timed("whole frame", loop._frame)
timed("talker decode pass (28 layers)", talker_pass)
timed("predictor decode pass (5 layers)", pred_decode)
timed("predictor prefill pass (2 tok/seq)", pred_prefill)
timed("lm_head + noise + sampler", head_sample)
lay = T.model.layers[3]
def one_layer():
    nat.decode_layout(loop.seq_lens, loop.t_tables, loop.page, 1, 0, loop.t_pos, loop.t_slots, loop.t_ctx)
    set_context(False, slot_mapping=loop.t_slots, context_lens=loop.t_ctx, block_tables=loop.t_tables,
                max_context_len=loop.width * loop.page)
    r = loop.embeds.clone()
    for _ in range(8):
        r, q = lay.forward_decode(loop.t_pos, r, None, True)
    reset_context()
timed("talker layer x8 (+layout, clone)", one_layer)
