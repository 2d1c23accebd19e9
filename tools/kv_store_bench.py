import sys; sys.path.insert(0, '.')
import torch, bench_kernels as BK
for n in (64, 4096, 16384):
    ms, alg = BK.kv_store(n, "cuda:0")
    print(f"store_kvcache {n} tokens: {ms*1e3:.2f} us {alg/ms/1e6:.0f} GB/s")
try:
    ms, alg = BK.kv_store(4096, "cuda:0", impl="triton"); print(f"triton 4096: {ms*1e3:.2f} us {alg/ms/1e6:.0f} GB/s")
except Exception as e: print("triton", e)
