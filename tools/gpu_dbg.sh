#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python - <<'PY'
import torch, time
from plaincv_b200 import ops
n = 405_000_000
p = torch.randn(n, device="cuda") * 0.02; g = torch.randn(n, device="cuda") * 0.01
m = torch.zeros(n, device="cuda"); v = torch.zeros(n, device="cuda"); pb = torch.zeros(n, dtype=torch.bfloat16, device="cuda")
partials = torch.zeros(592, device="cuda"); norm = torch.zeros(1, device="cuda")
for name, fn, bytes_ in (("sumsq", lambda: ops.sumsq_partials(g, partials), 4 * n),
                         ("adamw", lambda: ops.adamw_step(p, g, m, v, pb, partials, norm, 1.0, 3e-3, 0.9, 0.999, 1e-8, 0.1, 1), 30 * n)):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): fn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(f"{name}: {ms*1e3:.1f} us  {bytes_/ms/1e6:.0f} GB/s")
PY
