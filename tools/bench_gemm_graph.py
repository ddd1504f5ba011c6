#!/usr/bin/env python
"""Ours beside cuBLAS (torch.matmul) at the 420M GEMM shapes WITHOUT host effects: R launches of one GEMM captured in
a CUDA graph (rotating over enough operand sets to exceed L2, as consecutive layers do in a step), replayed and timed
with CUDA events; per-launch microseconds.  Measurement only — nothing on the product path calls cuBLAS."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from plaincv_b200 import ops  # noqa: E402
from plaincv_b200.ops import gemm, PCV_EPI_BF16, PCV_EPI_F32_ATOMIC  # noqa: E402

dev, bf = "cuda", torch.bfloat16
M, d, hp, Vp = 16384, 1024, 2752, 50304
R = int(os.environ.get("PCV_GG_R", 24))
SETS = int(os.environ.get("PCV_GG_SETS", 4))


def graph_time(fn, reps=3):
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


rows = []
shapes = (("qkv", d, 3 * d), ("out", d, d), ("gu", d, 2 * hp), ("down", hp, d))
if os.environ.get("PCV_GG_LM", "0") == "1":
    shapes = shapes + (("lm", d, Vp),)
if os.environ.get("PCV_GG_ONLY"):
    shapes = tuple(sh for sh in (shapes + (("lm", d, Vp),)) if sh[0] in os.environ["PCV_GG_ONLY"].split(","))
    shapes = tuple(dict.fromkeys(shapes))
NOCUBLAS = os.environ.get("PCV_GG_NOCUBLAS", "0") == "1"
for name, K, N in shapes:
    sets = SETS if name != "lm" else 1
    x = [torch.randn(M, K, device=dev).to(bf) for _ in range(sets)]
    w = [(torch.randn(K, N, device=dev) * 0.02).to(bf) for _ in range(sets)]
    dy = [torch.randn(M, N, device=dev).to(bf) for _ in range(sets)]
    y = [torch.empty(M, N, device=dev, dtype=bf) for _ in range(sets)]
    dx = [torch.empty(M, K, device=dev, dtype=bf) for _ in range(sets)]
    dwb = [torch.empty(K, N, device=dev, dtype=bf) for _ in range(sets)]
    dwf = [torch.zeros(K, N, device=dev) for _ in range(sets)]
    fl = 2.0 * M * K * N
    for op, ours, lib in (
        ("fwd", lambda i: gemm(x[i % sets], w[i % sets], y[i % sets], M, N, K, K, N, N, False, True, PCV_EPI_BF16),
         lambda i: torch.matmul(x[i % sets], w[i % sets], out=y[i % sets])),
        ("dgrad", lambda i: gemm(dy[i % sets], w[i % sets], dx[i % sets], M, K, N, N, N, K, False, False, PCV_EPI_BF16, 1, 1.0),
         lambda i: torch.matmul(dy[i % sets], w[i % sets].t(), out=dx[i % sets])),
        ("wgrad", lambda i: gemm(x[i % sets], dy[i % sets], dwf[i % sets], K, N, M, K, N, N, True, True, PCV_EPI_F32_ATOMIC, 0, 1.0),
         lambda i: torch.matmul(x[i % sets].t(), dy[i % sets], out=dwb[i % sets])),
    ):
        t_o = graph_time(ours)
        t_l = t_o if NOCUBLAS else graph_time(lib)
        rows.append((f"{op}_{name}", t_o, t_l))
        print(f"{op+'_'+name:14s} ours {t_o:8.1f} us {fl/t_o/1e6:8.1f} TFLOP/s | cublas {t_l:8.1f} us {fl/t_l/1e6:8.1f} TFLOP/s | ours/cublas time {t_o/t_l:5.3f}", flush=True)
    del x, w, dy, y, dx, dwb, dwf
    torch.cuda.empty_cache()
