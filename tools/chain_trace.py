"""In-kernel timeline of the persistent GEMM chain (debug): `chain` launches in one CUDA graph; the first epilogue
thread of every CTA stamps %globaltimer at 7 points per op.   python tools/chain_trace.py M H INTER NQ [launches]"""
import ctypes, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nano_qwen3tts_vllm_b200._native as nat

m, h, inter, nq = map(int, sys.argv[1:5])
chain = int(sys.argv[5]) if len(sys.argv) > 5 else 3
lib = nat.lib()
lib.b2tts_debug_gemm_trace.argtypes = [ctypes.c_void_p, ctypes.c_int]
DEV, BF16 = "cuda", torch.bfloat16
g = torch.Generator(device=DEV).manual_seed(1)
mk = lambda n, k: (torch.randn(n, k, device=DEV, generator=g) * 0.03).to(BF16)
L = [dict(o=nat.gemm_pack_weight(mk(h, nq // 2)), gu=nat.gemm_pack_weight(mk(2 * inter, h), nat.PACK_GATE_UP),
          down=nat.gemm_pack_weight(mk(h, inter)), qkv=nat.gemm_pack_weight(mk(nq, h))) for _ in range(2 * chain)]
ln = torch.ones(h, device=DEV, dtype=BF16)
attn = torch.randn(m, nq // 2, device=DEV, dtype=BF16)
r = torch.randn(m, h, device=DEV, dtype=BF16)
act = torch.empty(m, inter, device=DEV, dtype=BF16); qkv = torch.empty(m, nq, device=DEV, dtype=BF16)
s1 = torch.empty(m, h // 128, device=DEV, dtype=torch.float32); s2 = torch.empty_like(s1)
ws = nat.gemm_chain_workspace(m, [(h, nq // 2), (2 * inter, h), (h, inter), (nq, h)], DEV)


def run(ws_list):
    for w in ws_list:
        nat.gemm_chain([
            dict(x=attn, packed=w["o"], n=h, residual=r, out=r, epilogue=nat.GEMM_RESIDUAL, ssq_out=s1),
            dict(x=r, packed=w["gu"], n=2 * inter, out=act, epilogue=nat.GEMM_SILU_MUL, norm_weight=ln, ssq_in=s1),
            dict(x=act, packed=w["down"], n=h, residual=r, out=r, epilogue=nat.GEMM_RESIDUAL, ssq_out=s2),
            dict(x=r, packed=w["qkv"], n=nq, out=qkv, norm_weight=ln, ssq_in=s2)], m, ws)


run(L[:chain]); torch.cuda.synchronize()
trace = torch.zeros(chain * 4096 * 8, dtype=torch.int64, device=DEV)
lib.b2tts_debug_gemm_trace(trace.data_ptr(), chain)
gr = torch.cuda.CUDAGraph()
with torch.cuda.graph(gr):
    run(L[chain:])
lib.b2tts_debug_gemm_trace(None, 0)
gr.replay(); torch.cuda.synchronize()
trace.zero_()
gr.replay(); torch.cuda.synchronize()
t = trace.view(chain, 4096 * 8)[:, :148 * 4 * 8].view(chain, 148, 4, 8).cpu().double()
base = t[0, :, 0, 0][t[0, :, 0, 0] > 0].min()
names = ["entry", "dep ok", "rstd", "x landed", "drained", "barrier A", "reduced"]
ops = ["o_proj", "gate_up", "down", "qkv"]
print(f"M={m} H={h} I={inter} QKV={nq}; us since first CTA entry of launch 0; min/median/max over CTAs")
for l in range(chain):
    for i in range(4):
        tt = (t[l, :, i] - base) / 1e3
        cols = []
        for k in range(1 if i else 0, 7):
            v = tt[:, k][t[l, :, i, k] > 0]
            cols.append(f"{names[k]:>9s} {v.min():6.1f}/{v.median():6.1f}/{v.max():6.1f}" if v.numel() else f"{names[k]:>9s}   -")
        raw = trace.view(chain, 4096 * 8)[l, :148 * 4 * 8].view(148, 4, 8)[:, i, 7].cpu()
        ww, wx = (raw >> 32).double() / 1.965e3, (raw & 0xffffffff).double() / 1.965e3
        print(f"L{l} {ops[i]:8s}", "  ".join(cols), f"  MMA waits us: W {ww.median():.1f}/{ww.max():.1f}  x {wx.median():.1f}/{wx.max():.1f}")

kb = trace.view(chain, 4096 * 8)[chain - 1, 8192:8192 + 5 * 64].view(5, 64).cpu().double()
t0 = kb[kb > 0].min()
print("CTA 0, gate_up, per k-block (us since first stamp): x TMA issued | x landed (hop) | hop done | MMA has W | MMA has x")
for idx in range(64):
    if kb[0, idx] == 0:
        break
    print(f"  kb {idx:2d}: " + "  ".join(f"{(kb[e, idx] - t0) / 1965:7.2f}" for e in range(5)))
