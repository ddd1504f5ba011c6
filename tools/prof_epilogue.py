#!/usr/bin/env python
"""Per-phase cycle counts of the CTA-pair GEMM epilogue warps (pcv_gemm_set_profile) at the 420M MLP shapes.
GPU box only; needs a measurement build: make -C plaincv_b200/csrc clean all XFLAGS=-DPCV_MEASURE.
usage: python tools/prof_epilogue.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from plaincv_b200 import _lib, ops  # noqa: E402

dev = "cuda"
bf = torch.bfloat16
M, d, hp = 16384, 1024, 2752
lib = _lib.load()
buf = torch.zeros(148 * 8 * 8, dtype=torch.int32, device=dev)
xg = torch.randn(M, d, device=dev).to(bf)
wgu = (torch.randn(d, 2 * hp, device=dev) * 0.02).to(bf)
wdn = (torch.randn(hp, d, device=dev) * 0.02).to(bf)
gu_s, _ = ops.linear_fwd_swiglu(xg, wgu)
xh = torch.randn(M, hp, device=dev).to(bf)
cases = {
    "fwd gu+swiglu (19 tiles/pair)": lambda: ops.linear_fwd_swiglu(xg, wgu),
    "fwd gu plain (19 tiles/pair)": lambda: ops.linear_fwd(xg, wgu),
    "dgrad down+dswiglu (9.5 tiles/pair)": lambda: ops.linear_dgrad_dswiglu(xg, wdn, gu_s),
    "dgrad down plain (3.5 tiles/pair)": lambda: ops.linear_dgrad(xh[:, :d].contiguous(), wdn),
}
for name, fn in cases.items():
    fn()
    torch.cuda.synchronize()
    buf.zero_()
    import ctypes
    lib.pcv_gemm_set_profile.argtypes = [ctypes.c_void_p]  # -DPCV_MEASURE builds only
    lib.pcv_gemm_set_profile(buf.data_ptr())
    fn()
    torch.cuda.synchronize()
    lib.pcv_gemm_set_profile(None)
    t = buf.view(148, 8, 8).float()
    m = t.mean(dim=(0, 1))
    print(f"{name:38s} mean cycles/warp: wait_acc {m[0]:9.0f}  wait_operands {m[1]:9.0f}  math {m[2]:9.0f}  "
          f"stores {m[3]:9.0f}  total {m[4]:9.0f}")
