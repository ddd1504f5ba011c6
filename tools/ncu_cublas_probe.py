#!/usr/bin/env python
"""cuBLAS (torch.matmul) at the 420M GEMM shapes, one launch each, for an `ncu --set full` capture: which kernel,
grid / cluster / shared memory it picks, and its tensor-pipe utilisation, clocks and traffic beside ours
(profiles/r02_ncu_cublas_probe.txt).  Measurement only — nothing on the product path calls cuBLAS."""
import torch

dev, bf = "cuda", torch.bfloat16
M, d, hp = 16384, 1024, 2752
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
for name, K, N in (("qkv", d, 3 * d), ("out", d, d), ("gu", d, 2 * hp), ("down", hp, d)):
    x = torch.randn(M, K, device=dev).to(bf)
    w = (torch.randn(K, N, device=dev) * 0.02).to(bf)
    dy = torch.randn(M, N, device=dev).to(bf)
    y = torch.empty(M, N, device=dev, dtype=bf)
    dx = torch.empty(M, K, device=dev, dtype=bf)
    dw = torch.empty(K, N, device=dev, dtype=bf)
    for _ in range(2):
        flush.zero_()
        torch.matmul(x, w, out=y)
        flush.zero_()
        torch.matmul(dy, w.t(), out=dx)
        flush.zero_()
        torch.matmul(x.t(), dy, out=dw)
    torch.cuda.synchronize()
print("cublas probe done")
