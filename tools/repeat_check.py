#!/usr/bin/env python
"""Debug: run-to-run differences of the gradients at C3 shapes (2 layers), per parameter slot.  GPU box only."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from plaincv_b200.batch import synthetic_batch
from plaincv_b200.engine import Engine
from plaincv_b200.model import ModelConfig, Transformer

cfg = ModelConfig(vocab_size=50280, seq_len=2048, dim=1024, expand=8 / 3, n_layers=2, n_heads=16)
model = Transformer(cfg, device="cuda", seed=0)
eng = Engine(model, seq_len=2048)
batch = synthetic_batch(8, 2048, 50280, seed=1234)
L = model.layout
runs = []
for k in range(4):
    eng.compute_grads(batch)
    torch.cuda.synchronize()
    runs.append(model.flat_grad.clone())
names = ["embed", "layers_0.w_qkv", "layers_0.w_out", "layers_0.w_gu", "layers_0.w_down", "layers_1.w_qkv", "layers_1.w_out",
         "layers_1.w_gu", "layers_1.w_down", "w_lm"]
def rel(a, b):
    return float((a.double() - b.double()).norm() / (b.double().norm() + 1e-30))
print("whole buffer:", [f"{rel(runs[k], runs[0]):.2e}" for k in range(1, 4)])
for n in names:
    try:
        v = [L.slot_view(r, n) for r in runs]
    except Exception as e:
        print(n, "n/a", e)
        continue
    print(f"{n:18s}", [f"{rel(v[k], v[0]):.2e}" for k in range(1, 4)])
