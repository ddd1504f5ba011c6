#!/usr/bin/env python
"""RMSNorm forward / backward timed as 24 graph-captured launches over rotating buffers (no host effects, working set
beyond L2): per-launch microseconds and GB/s of the algorithmic bytes.  PCV_LIB_PATH selects an A/B build."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from plaincv_b200 import ops  # noqa: E402

dev, bf = "cuda", torch.bfloat16
M, d, R, SETS = 16384, 1024, 24, 6


def graph_time(fn, reps=5):
    fn(0)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(R):
            fn(i)
    g.replay()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / R)
    return best * 1e3


x = [torch.randn(M, d, device=dev).to(bf) for _ in range(SETS)]
dl = [torch.randn(M, d, device=dev).to(bf) for _ in range(SETS)]
scale = torch.ones(d, device=dev)
dscale = torch.zeros(d, device=dev)
outs = [ops.rmsnorm_residual_fwd(x[i], dl[i], scale, 1e-6) for i in range(SETS)]
t = graph_time(lambda i: ops.rmsnorm_residual_fwd(x[i % SETS], dl[i % SETS], scale, 1e-6))
print(f"rmsnorm_res_fwd {t:7.1f} us {4 * M * d * 2 / t / 1e3:8.1f} GB/s")
t = graph_time(lambda i: ops.rmsnorm_residual_bwd(x[i % SETS], dl[i % SETS], outs[i % SETS][0], scale, outs[i % SETS][2], dscale))
print(f"rmsnorm_res_bwd {t:7.1f} us {4 * M * d * 2 / t / 1e3:8.1f} GB/s")
