#!/usr/bin/env python
"""GGN matvec at the 420M (C3) shapes: CUDA-event time of one J^T H J v product through
plaincv_b200.curvature.make_ggn_matvec_fn_lm, the tangent kernels alone (achieved TFLOP/s or GB/s against the
measured peaks), and the training step's forward+backward beside it.  GPU box only; prints one JSON line.
usage: python tools/bench_curvature.py [n_layers]   (default 24 = C3)"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from plaincv_b200 import ops  # noqa: E402
from plaincv_b200.curvature import TangentBuffers, jvp_logits, make_ggn_matvec_fn_lm  # noqa: E402
from plaincv_b200.model import ModelConfig, Transformer  # noqa: E402

dev = "cuda"
bf = torch.bfloat16
L = int(sys.argv[1]) if len(sys.argv) > 1 else 24
B, T, NH, d, V = 8, 2048, 16, 1024, 50280
M = B * T
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
peaks = {"hbm_gbs": 6548.8, "bf16_tflops_sustained": 1381.0}
p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
if os.path.exists(p):
    peaks.update(json.load(open(p)))


def timeit(fn, iters=5, warm=2):
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


out = {"workload": f"GGN matvec, 420M LM widths, {L} layers, batch {B} x {T}", "kernels": {}}

# ---- tangent kernels alone
qkv = torch.randn(M, 3 * d, device=dev).to(bf)
tqkv = torch.randn(M, 3 * d, device=dev).to(bf)
o, lse = ops.attn_fwd(qkv, None, B, T, NH)
ms = timeit(lambda: ops.attn_jvp(qkv, tqkv, None, o, lse, B, T, NH))
fl = 5 * 2.0 * B * NH * T * T * 64 / 2  # five 128x128x64-class MMAs per visible (query, key) tile pair, causal half
out["kernels"]["attn_jvp"] = {"us": ms * 1e3, "tflops": fl / ms / 1e9, "frac_of_bf16_sustained": fl / ms / 1e9 / peaks["bf16_tflops_sustained"]}
ms_f = timeit(lambda: ops.attn_fwd(qkv, None, B, T, NH))
out["kernels"]["attn_fwd"] = {"us": ms_f * 1e3, "tflops": 2 * 2.0 * B * NH * T * T * 64 / 2 / ms_f / 1e9}

x = torch.randn(M, d, device=dev).to(bf)
tx = torch.randn(M, d, device=dev).to(bf)
td = torch.randn(M, d, device=dev).to(bf)
scale = torch.ones(d, device=dev)
res_out, y, rstd = ops.rmsnorm_residual_fwd(x, td, scale, 1e-6)
ms = timeit(lambda: ops.rmsnorm_residual_jvp(res_out, rstd, scale, scale, tx, td))
by = 5.0 * M * d * 2  # read x, tres, tdelta; write tres', ty
out["kernels"]["rmsnorm_residual_jvp"] = {"us": ms * 1e3, "gbs": by / ms / 1e6, "frac_of_hbm": by / ms / 1e6 / peaks["hbm_gbs"]}

hp = 2752
gu = torch.randn(M, 2 * hp, device=dev).to(bf)
tgu = torch.randn(M, 2 * hp, device=dev).to(bf)
ms = timeit(lambda: ops.swiglu_jvp(gu, tgu))
by = 5.0 * M * hp * 2
out["kernels"]["swiglu_jvp"] = {"us": ms * 1e3, "gbs": by / ms / 1e6, "frac_of_hbm": by / ms / 1e6 / peaks["hbm_gbs"]}

Vp = 50304
logits = (torch.randn(M, Vp, device=dev) * 2).to(bf)
tl = torch.randn(M, Vp, device=dev).to(bf)
ms = timeit(lambda: ops.ce_hessian_vec(logits, tl, V, 1.0 / M, out=tl))
by = 3.0 * M * Vp * 2  # read logits, read v, write hv (the re-reads of a row hit L1/L2)
out["kernels"]["ce_hessian_vec"] = {"us": ms * 1e3, "gbs": by / ms / 1e6, "frac_of_hbm": by / ms / 1e6 / peaks["hbm_gbs"]}
del logits, tl, gu, tgu, qkv, tqkv

# ---- the product
cfg = ModelConfig(vocab_size=V, seq_len=T, dim=d, expand=8 / 3, n_layers=L, n_heads=NH)
model = Transformer(cfg, device=dev, seed=0)
ids = torch.randint(0, V, (B, T), generator=torch.Generator().manual_seed(0)).to(dev)
vec = torch.randn(model.layout.total, device=dev) * 0.02
matvec = make_ggn_matvec_fn_lm(model, (ids, ids, None))
ms_mv = timeit(lambda: matvec(None, vec), iters=3, warm=1)
tb = TangentBuffers(model)
tb.load(vec)
ms_jvp = timeit(lambda: jvp_logits(model, ids, tb), iters=3, warm=1)


def fwd_bwd():
    model.flat_grad.zero_()
    loss, _ = model.loss_and_acc(ids, ids, unit_grad=True)
    loss.backward()


ms_fb = timeit(fwd_bwd, iters=3, warm=1)
hid = int(8 / 3 * d)
gemm_fl_tok = 2.0 * (L * (4 * d * d + 3 * d * hid) + d * V)  # one forward's Dense flops per token
# executed flops: one primal forward (1x), two GEMMs per Dense tangent (2x), dgrad + wgrad (2x); attention in units of
# one causal-half 128x128x64-class MMA sweep: forward 2, tangent 5, backward 5
model_fl = 5 * gemm_fl_tok * M + L * (2 + 5 + 5) * 2.0 * B * NH * T * T * 64 / 2
out.update({"matvec_ms": ms_mv, "jvp_logits_ms": ms_jvp, "train_fwd_bwd_ms": ms_fb,
            "matvec_over_fwd_bwd": ms_mv / ms_fb, "matvec_tflops": model_fl / ms_mv / 1e9,
            "frac_of_bf16_sustained": model_fl / ms_mv / 1e9 / peaks["bf16_tflops_sustained"],
            "tokens_per_s": M / (ms_mv / 1e3)})
print(json.dumps(out))
