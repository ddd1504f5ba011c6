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

def cta_report(tag, n_cta):
    cbuf = (ctypes.c_longlong * (4096 * 4))()
    lib.pcv_debug_read_cta_trace.argtypes = [ctypes.c_void_p, ctypes.c_int]
    lib.pcv_debug_read_cta_trace(cbuf, 4096 * 4)
    c = np.array(cbuf[:], dtype=np.int64).reshape(4096, 4)[:n_cta]
    c = c[(c[:, 0] > 0) & (c[:, 1] > c[:, 0])]  # persistent kernels: CTAs whose launch was cancelled never ran
    t_begin, t_end = c[:, 0].min(), c[:, 1].max()
    print(f"--- {tag} CTA occupancy: kernel span {(t_end - t_begin) / 1e3:.1f} us, {len(c)} CTAs on {len(set(c[:, 2]))} SMs")
    busy = (c[:, 1] - c[:, 0]).sum() / 1e3
    print(f"sum of CTA lifetimes {busy:.0f} us = {busy / 148:.1f} us per SM; mean lifetime (us) by number of key/query tiles visited:")
    for n in sorted(set(c[:, 3])):
        sel = c[c[:, 3] == n]
        print(f"  n={n:2d}: {((sel[:, 1] - sel[:, 0]).mean()) / 1e3:7.2f}", end="")
    print()
    gaps, last_end = [], []
    for sm in sorted(set(c[:, 2])):
        sel = c[c[:, 2] == sm]
        sel = sel[np.argsort(sel[:, 0])]
        gaps += list((sel[1:, 0] - sel[:-1, 1]) / 1e3)
        last_end.append((sel[-1, 1] - t_begin) / 1e3)
    gaps = np.array(gaps)
    print(f"gap between consecutive CTAs on an SM: mean {gaps.mean():.2f} us, median {np.median(gaps):.2f}, max {gaps.max():.2f}, total per SM {gaps.sum() / 148:.1f} us")
    le = np.array(last_end)
    print(f"SM finish times: min {le.min():.1f} median {np.median(le):.1f} max {le.max():.1f} us")

buf = (ctypes.c_longlong * 8192)()
lib.pcv_debug_read_trace.argtypes = [ctypes.c_void_p, ctypes.c_int]
print("rc", lib.pcv_debug_read_trace(buf, 8192))
a = np.array(buf[:], dtype=np.int64).reshape(8, 1024)
t0 = a[0, 0]
print("--- fwd CTA (0,0,0) = heaviest query tile, softmax warpgroup: per 64-key half tile [wait_S, pass1, pass2+arrive, loop] cycles; abs start")
s = a[0]
n = 2 * (T // 128)
for i in range(n):
    x = s[i * 4:i * 4 + 5]
    print(i, x[1] - x[0], x[2] - x[1], x[3] - x[2], (x[4] - x[3]) if i + 1 < n else -1, "  @", x[0] - t0)
m = a[4]
print("--- MMA waits for P: [wait cycles] @abs")
print([(int(m[i * 2 + 1] - m[i * 2]), int(m[i * 2] - t0)) for i in range(n)])

ld = a[6]
print("--- fwd loader (thread 256): per K tile [wait empty slot, load+rope+store] @abs")
print([(int(ld[i * 3 + 1] - ld[i * 3]), int(ld[i * 3 + 2] - ld[i * 3 + 1]), int(ld[i * 3] - t0)) for i in range(16)])

cta_report('fwd', B * NH * (T // 128))

# ---- backward
dout = torch.randn(B * T, d, device="cuda").to(torch.bfloat16)
ops.attn_bwd(qkv, cs, None, out, dout, lse, B, T, NH)
torch.cuda.synchronize()
lib.pcv_debug_clear_cta_trace()  # persistent kernel: CTAs whose launch is cancelled leave no record
ops.attn_bwd(qkv, cs, None, out, dout, lse, B, T, NH)
torch.cuda.synchronize()
lib.pcv_debug_read_trace(buf, 8192)
a = np.array(buf[:], dtype=np.int64).reshape(8, 1024)
t0 = a[0, 0]
for hf in range(2):
    print(f"--- bwd compute half{hf} (key tile 0: 16 q tiles): [wait_SdP, ld+math, wait dQfull(it-1), store P/dS+arrive] @abs")
    s = a[hf]
    for i in range(16):
        x = s[i * 4:i * 4 + 5]
        print(i, x[1] - x[0], x[2] - x[1], x[3] - x[2], (x[4] - x[3]) if i < 15 else -1, " @", x[0] - t0)
print("--- bwd compute half0 store phase: [sts issue, tmem_st+wait, fences, arrive, to next tile start]")
for i in range(15):
    x = a[2][i * 4:i * 4 + 4]
    print(i, int(x[0] - a[0][i * 4 + 3]), int(x[1] - x[0]), int(x[2] - x[1]), int(x[3] - x[2]), int(a[0][(i + 1) * 4] - x[3]))
m = a[4]
m5 = a[5]
print("--- bwd MMA per tile: [wait QdO+SdPfree, issue SdP(it+1), wait PdSfull(it), wait dQempty, issue dV/dK/dQ] @abs")
for i in range(16):
    print(i, int(m5[i * 4] - m[i * 2]), int(m5[i * 4 + 1] - m5[i * 4]), int(m[i * 2 + 1] - m5[i * 4 + 1]) if i < 15 else int(m[i*2+1]-m[i*2]),
          int(m5[i * 4 + 2] - m[i * 2 + 1]), int(m5[i * 4 + 3] - m5[i * 4 + 2]), " @", int(m[i * 2] - t0))

e = a[6]
print("--- bwd CTA(0,0,0) ends (cycles from CTA start): first S/dP ready", int(a[0][1] - e[5]), " loop done", int(e[0] - e[5]),
      " cs loaded", int(e[7] - e[5]), " dK/dV accumulated", int(e[1] - e[5]), " in registers", int(e[6] - e[5]), " dK/dV written", int(e[2] - e[5]), " all warps done", int(e[3] - e[5]),
      " TMEM freed", int(e[4] - e[5]), " | dQ warpgroup: loop done", int(a[7][0] - e[5]), " reduce-adds complete", int(a[7][1] - e[5]))
cta_report('bwd', B * NH * (T // 128))
