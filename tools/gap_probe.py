#!/usr/bin/env python
"""Per-launch time of kernels issued back to back on one stream (CUDA events around 200 launches) — compared with
the kernel's own duration (ncu) this gives the launch-to-launch gap a programmatic dependent launch could hide."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from plaincv_b200 import ops
dev = "cuda"; bf = torch.bfloat16
M, d, hp = 16384, 1024, 2752
x = torch.randn(M, d, device=dev).to(bf); dl = torch.randn(M, d, device=dev).to(bf)
scale = torch.ones(d, device=dev)
wo = (torch.randn(d, d, device=dev) * 0.02).to(bf)
wgu = (torch.randn(d, 2 * hp, device=dev) * 0.02).to(bf)
qkv = torch.randn(M, 3 * d, device=dev).to(bf)
def timeit(fn, n=200):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
# raw launches through the C ABI with preallocated outputs (no allocator / autograd in the loop)
from plaincv_b200 import _lib
lib = _lib.load(); st = torch.cuda.current_stream().cuda_stream
ro = torch.empty_like(x); y = torch.empty_like(x); rstd = torch.empty(M, device=dev)
yo = torch.empty(M, d, device=dev, dtype=bf); ygu = torch.empty(M, 2 * hp, device=dev, dtype=bf)
print("rmsnorm fwd  %.1f us/launch" % timeit(lambda: lib.pcv_rmsnorm_residual_fwd(x.data_ptr(), dl.data_ptr(), scale.data_ptr(), ro.data_ptr(), y.data_ptr(), rstd.data_ptr(), M, d, 1e-6, st)))
print("gemm out     %.1f us/launch" % timeit(lambda: lib.pcv_gemm_bf16(x.data_ptr(), wo.data_ptr(), yo.data_ptr(), M, d, d, d, d, d, 0, 1, 0, 1, 1.0, st)))
print("gemm gu      %.1f us/launch" % timeit(lambda: lib.pcv_gemm_bf16(x.data_ptr(), wgu.data_ptr(), ygu.data_ptr(), M, 2 * hp, d, d, 2 * hp, 2 * hp, 0, 1, 0, 1, 1.0, st)))
out = torch.empty(M, d, device=dev, dtype=bf); lse = torch.empty(8 * 16 * 2048, device=dev)
print("attn fwd     %.1f us/launch" % timeit(lambda: lib.pcv_attn_fwd(qkv.data_ptr(), None, out.data_ptr(), lse.data_ptr(), 8, 2048, 16, st), 50))
# empty-ish kernel: launch cost floor
t = torch.zeros(8, device=dev, dtype=bf)
print("add_bf16(8)  %.1f us/launch" % timeit(lambda: lib.pcv_add_bf16(t.data_ptr(), t.data_ptr(), t.data_ptr(), 8, st)))
