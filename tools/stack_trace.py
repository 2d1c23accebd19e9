"""Timeline of the persistent dataflow decoder stack (csrc/decoder_stack.cu) at the code predictor's size, plus its
graph-replay time against the separate launches.  Usage: python tools/stack_trace.py [rows] [context]"""
import ctypes
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nano_qwen3tts_vllm_b200._native as nat  # noqa: E402
from nano_qwen3tts_vllm_b200.config import StackConfig  # noqa: E402
from nano_qwen3tts_vllm_b200.models import qwen3_tts_share as share  # noqa: E402
from nano_qwen3tts_vllm_b200.utils.context import set_context  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
CTX = int(sys.argv[2]) if len(sys.argv) > 2 else 10
BF16, PAGE, L = torch.bfloat16, 256, 5
lib = nat.lib()
lib.b2tts_debug_stack_trace.argtypes = [ctypes.c_void_p]

cfg = StackConfig(hidden_size=1024, intermediate_size=3072, num_hidden_layers=L, num_attention_heads=16,
                  num_key_value_heads=8, head_dim=128, vocab_size=2048)
torch.set_default_dtype(BF16)
stack = share.DecoderStack(cfg, 0).to("cuda")
torch.set_default_dtype(torch.float32)
share.random_init_(stack, 1)
g = torch.Generator(device="cuda").manual_seed(0)
lens = torch.full((B,), CTX, dtype=torch.int32, device="cuda")
bt = torch.arange(B, dtype=torch.int32, device="cuda").view(B, 1)
slots = (bt.view(-1) * PAGE + CTX - 1).to(torch.int32)
pos = (lens.long() - 1).contiguous()
ws_f = nat.fused_decode_attn_workspace(B, 16, 128, PAGE, "cuda")
for layer in stack.layers:
    layer.self_attn.attn.k_cache = torch.randn(B, PAGE, 8, 128, device="cuda", generator=g).to(BF16)
    layer.self_attn.attn.v_cache = torch.randn(B, PAGE, 8, 128, device="cuda", generator=g).to(BF16)
    layer.self_attn.attn.fused_workspace = ws_f
x = (torch.randn(B, 1024, device="cuda", generator=g) * 0.5).to(BF16)
set_context(False, slot_mapping=slots, context_lens=lens, block_tables=bt, max_context_len=PAGE)


def timed(fn, n=50):
    fn(); fn()
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        for _ in range(4):
            fn()
    gr.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        gr.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n / 4 * 1000


share.DATAFLOW_STACK = False
t_sep = timed(lambda: stack.run_layers(x, pos))
share.DATAFLOW_STACK = True
t_df = timed(lambda: stack.run_layers(x, pos))
print(f"rows {B} ctx {CTX}: separate launches {t_sep:.1f} us / pass, dataflow {t_df:.1f} us / pass (graph replay, 5 layers)")

G = nat.device_info()["sm_count"]
trace = torch.zeros(G * 8 * 32, dtype=torch.int64, device="cuda")
lib.b2tts_debug_stack_trace(trace.data_ptr())
stack.run_layers(x, pos)
torch.cuda.synchronize()
lib.b2tts_debug_stack_trace(None)
t = trace.cpu().numpy().astype(np.int64).reshape(G, 8, 32)
so, sd = stack.dataflow_plan()
nq, no, ng, nd = 32, 8 * so, 48, 8 * sd
roles = {"qkv": range(0, nq), "o": range(nq, nq + no), "gu": range(nq + no, nq + no + ng),
         "down": range(nq + no + ng, nq + no + ng + nd), "none": range(nq + no + ng + nd, G)}
t0 = t[:, 0, 13][t[:, 0, 13] > 0].min()
print(f"kernel span {(t[:, 0, 14].max() - t0) / 1000:.1f} us; CTA start spread {(t[:, 0, 13].max() - t0) / 1000:.1f} us")
names = ["top", "accfull", "pub1", "Tawait", "att_pro", "att_loop", "normpub", "att_start", "att_done", "x_dep", "x_issued",
         "mma_first", "mma_last", "start", "end", "att_ret", "att_bar", "att_red", "", "", "w0", "w1", "w2", "w3"]
ORDER = [0, 9, 11, 10, 12, 1, 2, 3, 6, 7, 4, 5, 15, 20, 21, 22, 23, 16, 17, 8]
for l in range(L):
    print(f"--- layer {l} (us since kernel start; median [min..max] over the role's CTAs)")
    for rn, rr in roles.items():
        idx = list(rr)
        if not idx:
            continue
        parts = []
        for sidx in ORDER:
            nm = names[sidx]
            v = t[idx, l, sidx]
            v = v[v > 0]
            if v.size:
                parts.append(f"{nm} {np.median(v - t0) / 1000:.1f}[{(v.min() - t0) / 1000:.1f}..{(v.max() - t0) / 1000:.1f}]")
        print(f"  {rn:5s} " + "  ".join(parts))
