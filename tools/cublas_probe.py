"""Which kernels does the library (cuBLAS through F.linear) pick for the decode projections?  Run under
ncu --metrics gpu__time_duration.sum,launch__grid_size,launch__block_size,launch__registers_per_thread,launch__shared_mem_per_block_dynamic,launch__cluster_size
to list them: python tools/cublas_probe.py"""
import torch
import torch.nn.functional as F

shapes = [(64, 4096, 2048), (64, 2048, 2048), (64, 12288, 2048), (64, 2048, 6144),          # talker qkv / o / gate_up / down
          (64, 4096, 1024), (64, 1024, 2048), (64, 6144, 1024), (64, 1024, 3072), (64, 2048, 1024),   # predictor + head
          (1, 4096, 1024), (8, 4096, 1024), (128, 4096, 1024)]
for m, n, k in shapes:
    x = torch.randn(m, k, device="cuda", dtype=torch.bfloat16)
    w = torch.randn(n, k, device="cuda", dtype=torch.bfloat16)
    for _ in range(3):
        y = F.linear(x, w)
    torch.cuda.synchronize()
    print(m, n, k, float(y.float().abs().mean()))
