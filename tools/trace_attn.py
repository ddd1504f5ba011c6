#!/usr/bin/env python
"""Debug: run attn_fwd once with the -DPCV_TRACE build and print the phase timeline of CTA (0,0,0)."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from plaincv_b200 import _lib
_lib.LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libpcv_trace.so")
from plaincv_b200 import ops
B, T, NH = 8, 2048, 16
d = NH * 64
qkv = torch.randn(B * T, 3 * d, device="cuda").to(torch.bfloat16)
cs = ops.rope_table(64, T, 500000.0, "cuda")
for _ in range(3):
    out, lse = ops.attn_fwd(qkv, None, B, T, NH)
torch.cuda.synchronize()
lib = _lib.load()
buf = (ctypes.c_longlong * 8192)()
lib.pcv_debug_read_trace.argtypes = [ctypes.c_void_p, ctypes.c_int]
print("rc", lib.pcv_debug_read_trace(buf, 8192))
a = np.array(buf[:], dtype=np.int64).reshape(8, 1024)
t0 = a[0, 0]
for w in range(2):
    print(f"--- softmax WG{w}: per tile [wait_S, pass1, pass2+arrive, fold+loop] cycles; abs start")
    s = a[w * 2]
    n = 17 if w == 1 else 16
    for i in range(n):
        x = s[i * 4:i * 4 + 5]
        print(i, x[1] - x[0], x[2] - x[1], x[3] - x[2], (x[4] - x[3]) if i + 1 < n else -1, "  @", x[0] - t0)
for w in range(2):
    m = a[4 + w]
    print(f"--- MMA waits for P of WG{w}: [wait cycles] @abs")
    print([(int(m[i * 2 + 1] - m[i * 2]), int(m[i * 2] - t0)) for i in range(16)])

ld = a[6]
print("--- fwd loader (thread 256): per K tile [wait empty slot, load+rope+store] @abs")
print([(int(ld[i * 3 + 1] - ld[i * 3]), int(ld[i * 3 + 2] - ld[i * 3 + 1]), int(ld[i * 3] - t0)) for i in range(16)])

# ---- backward
dout = torch.randn(B * T, d, device="cuda").to(torch.bfloat16)
for _ in range(2):
    ops.attn_bwd(qkv, cs, None, out, dout, lse, B, T, NH)
torch.cuda.synchronize()
lib.pcv_debug_read_trace(buf, 8192)
a = np.array(buf[:], dtype=np.int64).reshape(8, 1024)
t0 = a[0, 0]
for hf in range(2):
    print(f"--- bwd compute half{hf} (key tile 0: 16 q tiles): [wait_SdP, dQ-readout+red, compute+arrive, loop] @abs")
    s = a[hf]
    for i in range(16):
        x = s[i * 4:i * 4 + 5]
        print(i, x[1] - x[0], x[2] - x[1], x[3] - x[2], (x[4] - x[3]) if i < 15 else -1, " @", x[0] - t0)
m = a[4]
print("--- bwd MMA: (wait for PdS, then until next wait) ")
print([(int(m[i * 2 + 1] - m[i * 2]), int(m[i * 2 + 2] - m[i * 2 + 1]) if i < 15 else -1) for i in range(16)])
m5 = a[5]
print("--- bwd MMA detail per tile: [PdS->QdO ready, issue SdP, wait dQempty, issue dV/dK/dQ + commits]")
print([(int(m5[i * 4] - m[i * 2 + 1]), int(m5[i * 4 + 1] - m5[i * 4]), int(m5[i * 4 + 2] - m5[i * 4 + 1]), int(m5[i * 4 + 3] - m5[i * 4 + 2])) for i in range(15)])
