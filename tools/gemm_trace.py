"""In-kernel timeline of the weight-streaming GEMM (debug): a chain of GEMMs in one CUDA graph with
programmatic dependent launch; every CTA stamps %globaltimer at 8 points.
    python tools/gemm_trace.py M N K [splits] [chain]"""
import ctypes, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nano_qwen3tts_vllm_b200._native as nat

m, n, k = map(int, sys.argv[1:4])
splits = int(sys.argv[4]) if len(sys.argv) > 4 else 0
chain = int(sys.argv[5]) if len(sys.argv) > 5 else 8
lib = nat.lib()
lib.b2tts_debug_gemm_trace.argtypes = [ctypes.c_void_p, ctypes.c_int]
ws = [(torch.randn(n, k, device="cuda") * 0.05).to(torch.bfloat16) for _ in range(chain * 3)]
pk = [nat.gemm_pack_weight(w) for w in ws]
x = torch.randn(m, k, device="cuda").to(torch.bfloat16)
out = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
for i in range(chain):
    nat.gemm_bf16(x, pk[i], n, out=out, splits=splits, pdl=True)
torch.cuda.synchronize()
trace = torch.zeros(chain * 4096 * 8, dtype=torch.int64, device="cuda")
lib.b2tts_debug_gemm_trace(trace.data_ptr(), chain)
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    for i in range(chain):
        nat.gemm_bf16(x, pk[chain + i], n, out=out, splits=splits, pdl=True)
lib.b2tts_debug_gemm_trace(None, 0)
g.replay(); torch.cuda.synchronize()
trace.zero_()
g.replay(); torch.cuda.synchronize()
t = trace.view(chain, 4096, 8).cpu()
ncta = int((t[0, :, 0] != 0).sum())
names = ["entry", "setup done", "x: dep resolved", "x: first block staged", "accum complete", "staged", "exchange done", "end"]
base = int(t[0, :ncta, 0].min())
print(f"M={m} N={n} K={k} splits={splits} CTAs={ncta}; times in us relative to first CTA entry of launch 0")
for i in range(chain):
    tt = (t[i, :ncta].double() - base) / 1e3
    row = "  ".join(f"{names[j][:9]:>9s} {tt[:, j].min():5.1f}/{tt[:, j].median():5.1f}/{tt[:, j].max():5.1f}" for j in (0, 1, 2, 3, 4, 5, 6, 7))
    print(f"launch {i}: {row}")
