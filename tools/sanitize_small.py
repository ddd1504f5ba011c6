#!/usr/bin/env python
"""Every kernel of libplaincv_b200.so once, at small shapes, for compute-sanitizer (memcheck / racecheck / synccheck):
    compute-sanitizer --tool memcheck python tools/sanitize_small.py
Small on purpose (the tools slow kernels down 10-100x); the shapes still cross tile edges, use both GEMM tilings,
split-K, packed documents and several work items per persistent CTA."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from plaincv_b200 import _lib, ops  # noqa: E402
from plaincv_b200.batch import synthetic_batch  # noqa: E402
from plaincv_b200.engine import Engine  # noqa: E402
from plaincv_b200.model import ModelConfig, Transformer  # noqa: E402
from plaincv_b200.optim import AdamW  # noqa: E402

lib = _lib.load()
dev = "cuda"
bf = torch.bfloat16
torch.manual_seed(0)
done = []


def step(name, fn):
    fn()
    torch.cuda.synchronize()
    done.append(name)


# GEMMs: both tilings, every operand-major combination, ragged edges, split-K
for mode in (1, 2):
    lib.pcv_gemm_set_mode(mode)
    x = torch.randn(300, 264, device=dev).to(bf)
    w = (torch.randn(264, 520, device=dev) * 0.05).to(bf)
    dy = torch.randn(300, 520, device=dev).to(bf)
    dw = torch.zeros(264, 520, device=dev)
    step(f"gemm fwd mode {mode}", lambda: ops.linear_fwd(x, w))
    step(f"gemm dgrad mode {mode}", lambda: ops.linear_dgrad(dy, w))
    step(f"gemm wgrad mode {mode}", lambda: ops.linear_wgrad(x, dy, dw))
lib.pcv_gemm_set_mode(0)
B, T, NH = 2, 256, 2
d = NH * 64
cs = ops.rope_table(64, T, 500000.0, dev)
xq = torch.randn(B * T, d, device=dev).to(bf)
wq = (torch.randn(d, 3 * d, device=dev) * 0.05).to(bf)
step("gemm + rope", lambda: ops.linear_fwd_rope(xq, wq, cs, T, 2 * d))
hid = 344
wgu = (torch.randn(d, 2 * hid, device=dev) * 0.05).to(bf)
wdn = (torch.randn(hid, d, device=dev) * 0.05).to(bf)
gu_h = []
step("gemm + swiglu", lambda: gu_h.append(ops.linear_fwd_swiglu(xq, wgu)))
step("gemm + dswiglu", lambda: ops.linear_dgrad_dswiglu(xq, wdn, gu_h[0][0]))
# whole model (embedding, RMSNorm fwd/bwd, attention fwd/bwd with packed documents, fused head + loss, optimizer)
cfg = ModelConfig(vocab_size=1003, seq_len=256, dim=128, expand=8 / 3, n_layers=2, n_heads=2)
model = Transformer(cfg, device=dev, seed=0)
for use_doc in (False, True):
    eng = Engine(model, seq_len=256, use_doc_mask=use_doc)
    batch = synthetic_batch(3, 256, 1003, seed=1, doc_mean=60 if use_doc else None)
    step(f"model fwd+bwd doc={use_doc}", lambda: eng.compute_grads(batch))
opt = AdamW(model, lr=1e-3, weight_decay=0.1, grad_clip=1.0)
step("clip + adamw", lambda: opt.step())
# attention alone: several items per persistent backward CTA needs more CTAs than SMs: 5 x 8 heads x 512 / 128 = 160 items
Bq, Tq, NHq = 5, 512, 8
qkv = torch.randn(Bq * Tq, 3 * NHq * 64, device=dev).to(bf)
csq = ops.rope_table(64, Tq, 500000.0, dev)
o_l = []
step("attn fwd", lambda: o_l.append(ops.attn_fwd(qkv, None, Bq, Tq, NHq)))
dout = torch.randn(Bq * Tq, NHq * 64, device=dev).to(bf)
step("attn bwd (work stealing)", lambda: ops.attn_bwd(qkv, csq, None, o_l[0][0], dout, o_l[0][1], Bq, Tq, NHq))
# un-fused loss, stand-alone element-wise kernels, tangent kernels
logits = torch.randn(64, 1008, device=dev).to(bf)
labels = torch.randint(0, 1003, (64,), device=dev, dtype=torch.int32)
step("softmax_ce", lambda: ops.softmax_ce(logits, labels, 1003, 1.0 / 64, True))
gu = torch.randn(64, 2 * hid, device=dev).to(bf)
step("swiglu fwd/bwd/jvp", lambda: (ops.swiglu_fwd(gu), ops.swiglu_bwd(gu, gu[:, :hid].contiguous()), ops.swiglu_jvp(gu, gu)))
step("act fwd/bwd", lambda: (ops.act_fwd(gu, "mlp"), ops.act_bwd(gu, gu, "mlp_relu_sq")))
step("ce hessian-vec", lambda: ops.ce_hessian_vec(logits, logits.clone(), 1003, 1.0))
print("sanitize_small: ran", len(done), "steps:", ", ".join(done))
