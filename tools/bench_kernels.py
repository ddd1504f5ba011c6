#!/usr/bin/env python
"""Per-kernel CUDA-event timings at the 420M (C3) shapes; prints achieved TFLOP/s or GB/s.  GPU box only.
usage: python tools/bench_kernels.py [T] [which]   (which: all | attn | ew | gemm)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from plaincv_b200 import _lib  # noqa: E402
if os.environ.get("PCV_KB_LIB"):  # A/B builds of the library (tools/ab/*.so), measurement only
    _lib.LIB_PATH = os.path.abspath(os.environ["PCV_KB_LIB"])
from plaincv_b200 import ops  # noqa: E402

dev = "cuda"
bf = torch.bfloat16
B, T, NH, d, hp, Vp = 8, 2048, 16, 1024, int(os.environ.get("PCV_KB_HP", 2752)), int(os.environ.get("PCV_KB_VP", 50304))
if len(sys.argv) > 1:
    T = int(sys.argv[1])
    B = 16384 // T
which = sys.argv[2] if len(sys.argv) > 2 else "all"
M = B * T
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
res = {}


def timeit(fn, iters=int(os.environ.get('PCV_KB_ITERS', 10)), warm=int(os.environ.get('PCV_KB_WARM', 3))):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / iters


if which in ("all", "attn"):
    qkv = torch.randn(M, 3 * d, device=dev).to(bf)
    cos_sin = ops.rope_table(64, T, 500000.0, dev)
    out, lse = ops.attn_fwd(qkv, None, B, T, NH)
    dout = torch.randn(M, d, device=dev).to(bf)
    fl_fwd = 4.0 * B * NH * T * T * 64 / 2
    ms = timeit(lambda: ops.attn_fwd(qkv, None, B, T, NH))
    res["attn_fwd"] = (ms, fl_fwd / ms / 1e9, "TFLOP/s (causal-half flops)")
    ms = timeit(lambda: ops.attn_bwd(qkv, cos_sin, None, out, dout, lse, B, T, NH))
    res["attn_bwd"] = (ms, 2.5 * fl_fwd / ms / 1e9, "TFLOP/s (5 GEMMs, causal-half)")
    ds = torch.zeros(B, T, dtype=torch.int32)
    g = torch.Generator().manual_seed(0)
    for b in range(B):
        t = 0
        while t < T:
            n = int(max(1, min(T - t, torch.empty(1).exponential_(1 / 512.0, generator=g).item())))
            ds[b, t:t + n] = t
            t += n
    ds = ds.to(dev).view(-1)
    ms = timeit(lambda: ops.attn_fwd(qkv, ds, B, T, NH))
    res["attn_fwd_doc512"] = (ms, 0, "")
    ms = timeit(lambda: ops.attn_bwd(qkv, cos_sin, ds, out, dout, lse, B, T, NH))
    res["attn_bwd_doc512"] = (ms, 0, "")

if which in ("all", "ew"):
    x = torch.randn(M, d, device=dev).to(bf)
    dl = torch.randn(M, d, device=dev).to(bf)
    scale = torch.ones(d, device=dev)
    dscale = torch.zeros(d, device=dev)
    ro, y, rstd = ops.rmsnorm_residual_fwd(x, dl, scale, 1e-6)
    ms = timeit(lambda: ops.rmsnorm_residual_fwd(x, dl, scale, 1e-6))
    res["rmsnorm_res_fwd"] = (ms, 4 * M * d * 2 / ms / 1e6, "GB/s")
    ms = timeit(lambda: ops.rmsnorm_residual_bwd(x, dl, ro, scale, rstd, dscale))
    res["rmsnorm_res_bwd"] = (ms, 4 * M * d * 2 / ms / 1e6, "GB/s")
    gu = torch.randn(M, 2 * hp, device=dev).to(bf)
    dh = torch.randn(M, hp, device=dev).to(bf)
    ms = timeit(lambda: ops.swiglu_fwd(gu))
    res["swiglu_fwd"] = (ms, 3 * M * hp * 2 / ms / 1e6, "GB/s")
    ms = timeit(lambda: ops.swiglu_bwd(gu, dh))
    res["swiglu_bwd"] = (ms, 5 * M * hp * 2 / ms / 1e6, "GB/s")
    logits = (torch.randn(M, Vp, device=dev) * 2).to(bf)
    labels = torch.randint(0, Vp, (M,), device=dev, dtype=torch.int32)
    ms = timeit(lambda: ops.softmax_ce(logits, labels, Vp, 1.0 / M, True))
    res["softmax_ce(+grad)"] = (ms, 2 * M * Vp * 2 / ms / 1e6, "GB/s")
    del logits

if which == "ce":
    # LM head + loss, un-fused (pcv_gemm_bf16 + pcv_softmax_ce) against fused (pcv_gemm_bf16_ce + pcv_ce_from_stats)
    class _H:
        pass
    V = 50280
    head = _H()
    head.tied, head.vocab, head.v_pad = False, V, Vp
    head.w_lm = (torch.randn(d, Vp, device=dev) * 0.02).to(bf)
    hf = torch.randn(M, d, device=dev).to(bf)
    labels = torch.randint(0, V, (M,), device=dev, dtype=torch.int32)
    ms = timeit(lambda: ops.head_logits(hf, head))
    res["head_gemm"] = (ms, 2.0 * M * d * Vp / ms / 1e9, "TFLOP/s")
    ms = timeit(lambda: ops.head_logits_ce(hf, head))
    res["head_gemm+ce_partials"] = (ms, 2.0 * M * d * Vp / ms / 1e9, "TFLOP/s")
    logits, ce_ms = ops.head_logits_ce(hf, head)
    ms = timeit(lambda: ops.softmax_ce(logits, labels, V, 1.0 / M, True))
    res["softmax_ce(+grad)"] = (ms, 2 * M * Vp * 2 / ms / 1e6, "GB/s")
    logits, ce_ms = ops.head_logits_ce(hf, head)
    ms = timeit(lambda: ops.ce_from_stats(logits, labels, ce_ms, V, 1.0 / M))
    res["ce_from_stats(+grad)"] = (ms, 2 * M * Vp * 2 / ms / 1e6, "GB/s")
    # the same second half writing d(logits) to a separate buffer (is the in-place update what is slow?)
    from plaincv_b200 import _lib as _L
    lib = _L.load()
    out_buf = torch.empty_like(logits)
    rl = torch.empty(M, device=dev)
    rc_ = torch.empty(M, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    ms = timeit(lambda: lib.pcv_ce_from_stats(logits.data_ptr(), labels.data_ptr(), ce_ms.data_ptr(), ce_ms.shape[0],
                                              out_buf.data_ptr(), rl.data_ptr(), rc_.data_ptr(), M, V, Vp, 1.0 / M, st))
    res["ce_from_stats out-of-place"] = (ms, 2 * M * Vp * 2 / ms / 1e6, "GB/s")

if which == "fused":
    xg = torch.randn(M, d, device=dev).to(bf)
    wgu = (torch.randn(d, 2 * hp, device=dev) * 0.02).to(bf)
    wdn = (torch.randn(hp, d, device=dev) * 0.02).to(bf)
    dh = torch.randn(M, hp, device=dev).to(bf)
    for rep in range(2):
        gu_s, _ = ops.linear_fwd_swiglu(xg, wgu)
        ms = timeit(lambda: ops.linear_fwd_swiglu(xg, wgu))
        res[f"gemm_fwd_gu+swiglu#{rep}"] = (ms, 2.0 * M * d * 2 * hp / ms / 1e9, "TFLOP/s")
        ms = timeit(lambda: ops.linear_fwd(xg, wgu))
        res[f"gemm_fwd_gu#{rep}"] = (ms, 2.0 * M * d * 2 * hp / ms / 1e9, "TFLOP/s")
        ms = timeit(lambda: ops.swiglu_fwd(gu_s))
        res[f"swiglu_fwd#{rep}"] = (ms, 3 * M * hp * 2 / ms / 1e6, "GB/s")
        ms = timeit(lambda: ops.linear_dgrad_dswiglu(xg, wdn, gu_s))
        res[f"gemm_dgrad_down+dswiglu#{rep}"] = (ms, 2.0 * M * d * hp / ms / 1e9, "TFLOP/s")
        ms = timeit(lambda: ops.linear_dgrad(xg, wdn))
        res[f"gemm_dgrad_down#{rep}"] = (ms, 2.0 * M * d * hp / ms / 1e9, "TFLOP/s")
        ms = timeit(lambda: ops.swiglu_bwd(gu_s, dh))
        res[f"swiglu_bwd#{rep}"] = (ms, 5 * M * hp * 2 / ms / 1e6, "GB/s")

if which == "gemmstep":
    # every GEMM launch of one transformer block (forward, then backward) and of the LM head, once each, in the order a
    # training step issues them: the input of tools/gemm_traffic.py (DRAM bytes per launch from `ncu --set full`)
    xd = torch.randn(M, d, device=dev).to(bf)
    xh = torch.randn(M, hp, device=dev).to(bf)
    y3 = torch.randn(M, 3 * d, device=dev).to(bf)
    y2h = torch.randn(M, 2 * hp, device=dev).to(bf)
    wq = (torch.randn(d, 3 * d, device=dev) * 0.02).to(bf)
    wo = (torch.randn(d, d, device=dev) * 0.02).to(bf)
    wgu = (torch.randn(d, 2 * hp, device=dev) * 0.02).to(bf)
    wdn = (torch.randn(hp, d, device=dev) * 0.02).to(bf)
    wlm = (torch.randn(d, Vp, device=dev) * 0.02).to(bf)
    cs_t = ops.rope_table(64, T, 500000.0, dev)
    gq, go, ggu, gdn, glm = (torch.zeros_like(w, dtype=torch.float32) for w in (wq, wo, wgu, wdn, wlm))

    class _H:
        pass
    head_lm = _H()
    head_lm.tied, head_lm.vocab, head_lm.v_pad, head_lm.w_lm = False, 50280, Vp, wlm

    def layer_and_head():
        ops.linear_fwd_rope(xd, wq, cs_t, T, 2 * d)
        ops.linear_fwd(xd, wo)
        gu_s, _ = ops.linear_fwd_swiglu(xd, wgu)
        ops.linear_fwd(xh, wdn)
        ops.linear_dgrad_dswiglu(xd, wdn, gu_s)
        ops.linear_wgrad(xh, xd, gdn)
        ops.linear_dgrad(y2h, wgu)
        ops.linear_wgrad(xd, y2h, ggu)
        ops.linear_dgrad(xd, wo)
        ops.linear_wgrad(xd, xd, go)
        ops.linear_dgrad(y3, wq)
        ops.linear_wgrad(xd, y3, gq)
        logits, _ = ops.head_logits_ce(xd, head_lm)
        ops.linear_dgrad(logits, wlm)
        ops.linear_wgrad(xd, logits, glm)

    for _ in range(int(os.environ.get("PCV_KB_WARM", 1))):
        layer_and_head()
    torch.cuda.synchronize()
    flush.zero_()
    torch.cuda.synchronize()
    layer_and_head()
    torch.cuda.synchronize()
    print("gemmstep: 12 block GEMMs + 3 LM-head GEMMs launched")

if which in ("all", "gemm"):
    def gemm_set(name, K, N):
        xx = torch.randn(M, K, device=dev).to(bf)
        w = (torch.randn(K, N, device=dev) * 0.02).to(bf)
        dy = torch.randn(M, N, device=dev).to(bf)
        dw = torch.zeros(K, N, device=dev)
        fl = 2.0 * M * K * N
        ms = timeit(lambda: ops.linear_fwd(xx, w))
        res[f"gemm_fwd_{name}"] = (ms, fl / ms / 1e9, "TFLOP/s")
        ms = timeit(lambda: ops.linear_dgrad(dy, w))
        res[f"gemm_dgrad_{name}"] = (ms, fl / ms / 1e9, "TFLOP/s")
        ms = timeit(lambda: ops.linear_wgrad(xx, dy, dw))
        res[f"gemm_wgrad_{name}"] = (ms, fl / ms / 1e9, "TFLOP/s")

    def cublas_set(name, K, N):
        """library baseline beside each of ours (torch.matmul -> cuBLASLt), same shapes; not on the product path"""
        xx = torch.randn(M, K, device=dev).to(bf)
        w = (torch.randn(K, N, device=dev) * 0.02).to(bf)
        dy = torch.randn(M, N, device=dev).to(bf)
        fl = 2.0 * M * K * N
        y = torch.empty(M, N, device=dev, dtype=bf)
        dx = torch.empty(M, K, device=dev, dtype=bf)
        dw = torch.empty(K, N, device=dev, dtype=bf)
        ms = timeit(lambda: torch.matmul(xx, w, out=y))
        res[f"cublas_fwd_{name}"] = (ms, fl / ms / 1e9, "TFLOP/s")
        ms = timeit(lambda: torch.matmul(dy, w.t(), out=dx))
        res[f"cublas_dgrad_{name}"] = (ms, fl / ms / 1e9, "TFLOP/s")
        ms = timeit(lambda: torch.matmul(xx.t(), dy, out=dw))
        res[f"cublas_wgrad_{name}"] = (ms, fl / ms / 1e9, "TFLOP/s (bf16 out)")

    if os.environ.get("PCV_KB_CUBLAS", "1") == "1" and int(os.environ.get("PCV_KB_ITERS", 10)) > 1:
        for nm, kk, nn in (("qkv", d, 3 * d), ("out", d, d), ("gu", d, 2 * hp), ("down", hp, d), ("lm", d, Vp)):
            cublas_set(nm, kk, nn)
    xg = torch.randn(M, d, device=dev).to(bf)
    wgu = (torch.randn(d, 2 * hp, device=dev) * 0.02).to(bf)
    wdn = (torch.randn(hp, d, device=dev) * 0.02).to(bf)
    gu_s, _ = ops.linear_fwd_swiglu(xg, wgu)
    ms = timeit(lambda: ops.linear_fwd_swiglu(xg, wgu))
    res["gemm_fwd_gu+swiglu"] = (ms, 2.0 * M * d * 2 * hp / ms / 1e9, "TFLOP/s")
    ms = timeit(lambda: ops.linear_dgrad_dswiglu(xg, wdn, gu_s))
    res["gemm_dgrad_down+dswiglu"] = (ms, 2.0 * M * d * hp / ms / 1e9, "TFLOP/s")
    del xg, wgu, wdn, gu_s
    gemm_set("qkv", d, 3 * d)
    xx = torch.randn(M, d, device=dev).to(bf)
    wq = (torch.randn(d, 3 * d, device=dev) * 0.02).to(bf)
    cs_t = ops.rope_table(64, T, 500000.0, dev)
    ms = timeit(lambda: ops.linear_fwd_rope(xx, wq, cs_t, T, 2 * d))
    res["gemm_fwd_qkv+rope"] = (ms, 2.0 * M * d * 3 * d / ms / 1e9, "TFLOP/s")
    gemm_set("out", d, d)
    gemm_set("gu", d, 2 * hp)
    gemm_set("down", hp, d)
    gemm_set("lm", d, Vp)

for k, (ms, v, u) in res.items():
    print(f"{k:24s} {ms*1e3:9.1f} us   {v:9.1f} {u}")
