"""Parity at the widths the benchmark runs at.

(1) Whole model, loss + every parameter gradient against the fp32 CPU oracle, at the layer shapes of BASELINE.json's
    configs C2 / C3 / C4 / C5 with two layers (reference: models/LM/transformer.py:346-407,
    config/tr_420M_x8gpu.yaml:8-41, config/config_doc_mask.yaml:8-41) — the CTA-pair (cta_group::2) GEMM with its halved
    tail wave, library-chosen split-K and the fused RoPE / SwiGLU epilogues are the kernels that run here, and the test
    asserts that through pcv_gemm_pair_launch_count.
(2) The LM-head GEMMs at their real C3 shapes ([16384,1024] x [1024,50304] forward, dgrad with K = 50304, split-K
    wgrad with K = 16384 into fp32) against an fp32 torch.matmul on the GPU, full tensors.
(3) The cases of csrc/test_gemm.cu (every operand-major combination, ragged edges, split-K, several tiles per CTA)
    for both tilings against fp32 torch.matmul.

Tolerances are the whole-path ones of tests/test_model_gpu.py: loss rel. error <= 1e-3; per-parameter gradient
rel. L2 <= max(1e-2, 1.25 x the error of the oracle's own bf16-emulating mode), never above 3e-2 (the emulating pass is
only run when a leaf exceeds 1e-2)."""
import math

import numpy as np
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import lm_oracle as O  # noqa: E402  (test infrastructure: the checker)


def rel_l2(a, b):
    """||a - b|| / ||b|| in fp64; tensors that are both on the GPU are reduced there, a row block at a time."""
    a, b = a.detach(), b.detach()
    if a.is_cuda and b.is_cuda and a.dim() == 2:
        num = torch.zeros((), dtype=torch.float64, device=a.device)
        den = torch.zeros((), dtype=torch.float64, device=a.device)
        for r in range(0, a.shape[0], 2048):
            x, y = a[r:r + 2048].double(), b[r:r + 2048].double()
            num += ((x - y) ** 2).sum()
            den += (y ** 2).sum()
        return float(torch.sqrt(num) / (torch.sqrt(den) + 1e-30))
    a = a.double().cpu()
    b = b.double().cpu()
    return float((a - b).norm() / (b.norm() + 1e-30))


def _lib():
    from plaincv_b200 import _lib as L
    return L.load()


def _short_docs(B, T, n):
    """Every document n tokens long (the 64-token extreme of BASELINE.md §4); the +1 of the loader's T+1 contract is
    a one-token document that train_lm.py:_trim_last_token drops."""
    return [[n] * (T // n) + [1] for _ in range(B)]


# name -> (vocab, seq, d, heads, per-test batch, doc spec, gemm mode)
FULL_WIDTH = {
    # C3: config/tr_420M_x8gpu.yaml widths.  B = 3 (M = 6144), automatic tiling: all forward / dgrad GEMMs fill the 74
    # pairs, the N = 1024 ones (96 tiles, 22 in the last wave) run the halved 256 x 128 tail; split-K picked by the
    # library for the wgrads (the [1024, 1024] wgrad alone stays on single-CTA tiles at this M)
    "c3_420M_widths": dict(V=50280, T=2048, d=1024, nh=16, B=3, doc=None, mode=0),
    # the same shapes with the pair tiling forced for every launch, B = 2
    "c3_forced_pair": dict(V=50280, T=2048, d=1024, nh=16, B=2, doc=None, mode=2),
    # C2: GPT-2-small widths, d = 768 / 12 heads / hid 2048 / V = 50257 (not a multiple of 8: padded to 50304)
    # (M = 4096 here against 16384 in the benchmark: the pair tiling the benchmark gets is forced)
    "c2_gpt2_small_widths": dict(V=50257, T=1024, d=768, nh=12, B=4, doc=None, mode=2),
    # C4: the 420M widths at seq 4096
    "c4_seq4096": dict(V=50280, T=4096, d=1024, nh=16, B=1, doc=None, mode=2),
    # C5: config/config_doc_mask.yaml widths, packed documents (mean 512) ...
    "c5_docmask_mean512": dict(V=50280, T=2048, d=768, nh=12, B=2, doc=512, mode=2),
    # ... and the extreme of 64-token documents
    "c5_docmask_64tok": dict(V=50280, T=2048, d=768, nh=12, B=2, doc="short64", mode=2),
}


@pytest.mark.parametrize("name", list(FULL_WIDTH))
def test_full_width_parity(name):
    from plaincv_b200.engine import Engine
    from plaincv_b200.model import ModelConfig, Transformer
    c = FULL_WIDTH[name]
    V, T, d, nh, B = c["V"], c["T"], c["d"], c["nh"], c["B"]
    cfg_o = O.ModelConfig(vocab_size=V, seq_len=T, dim=d, expand=8 / 3, n_layers=2, n_heads=nh)
    cfg = ModelConfig(vocab_size=V, seq_len=T, dim=d, expand=8 / 3, n_layers=2, n_heads=nh)
    model = Transformer(cfg, device="cuda", seed=0)
    params = O.init_params(cfg_o, seed=0)
    model.load_flax_params(params)
    use_doc = c["doc"] is not None
    batch = O.synthetic_batch(B, T, V, seed=77, doc_mean=c["doc"] if isinstance(c["doc"], int) else None)
    if c["doc"] == "short64":
        batch["docs_lengths"] = _short_docs(B, T, 64)
    lib = _lib()
    lib.pcv_gemm_set_mode(c["mode"])
    try:
        eng = Engine(model, seq_len=T, use_doc_mask=use_doc)
        n0, p0 = lib.pcv_launch_count(), lib.pcv_gemm_pair_launch_count()
        loss, acc = eng.compute_grads(batch)
        torch.cuda.synchronize()
        pair_launches = lib.pcv_gemm_pair_launch_count() - p0
        launches = lib.pcv_launch_count() - n0
    finally:
        lib.pcv_gemm_set_mode(0)
    # 12 GEMMs per block (4 fwd, 4 dgrad, 4 wgrad) + 3 of the LM head
    n_gemm = 12 * cfg.n_layers + 3
    if c["mode"] == 2:
        assert pair_launches == n_gemm, (pair_launches, n_gemm, launches)
    else:
        assert n_gemm - 2 * cfg.n_layers <= pair_launches <= n_gemm, (pair_launches, n_gemm, launches)
    inputs, labels, mask = O.prepare_batch(batch, T, use_doc)
    g_ref, loss_ref, acc_ref = O.compute_grads(params, cfg_o, inputs, labels, mask, torch.float32)
    assert abs(float(loss) - float(loss_ref)) / float(loss_ref) < 1e-3, (float(loss), float(loss_ref))
    assert abs(float(acc) - float(acc_ref)) <= 2.0 / (B * T)
    grads = eng.grads()
    errs = {k: rel_l2(grads[k], g_ref[k]) for k in g_ref}
    worst = max(errs, key=errs.get)
    print(f"[{name}] loss {float(loss):.5f} oracle {float(loss_ref):.5f}; worst per-leaf gradient rel-L2 "
          f"{errs[worst]:.3e} ({worst}); pair-GEMM launches {pair_launches} of {launches} kernel launches")
    over = {k: e for k, e in errs.items() if e > 1e-2}
    if over:  # judge those leaves against what bf16 arithmetic itself costs (the oracle's bf16-emulating mode)
        g_emu, _, _ = O.compute_grads(params, cfg_o, inputs, labels, mask, torch.bfloat16)
        bad = {k: (e, rel_l2(g_emu[k], g_ref[k])) for k, e in over.items()
               if e > 3e-2 or e > 1.25 * rel_l2(g_emu[k], g_ref[k])}
        assert not bad, bad


# ------------------------------------------------------------------------------------------------ LM head, real shapes
def _fp32_matmul_chunks(a, b, rows=2048):
    """a @ b with fp32 operands and fp32 accumulation (TF32 off), a row block at a time."""
    out = torch.empty(a.shape[0], b.shape[1], dtype=torch.float32, device=a.device)
    for r in range(0, a.shape[0], rows):
        out[r:r + rows] = a[r:r + rows].float() @ b
    return out


@pytest.mark.parametrize("V", [50280, 50257])
def test_lm_head_gemms_at_full_shape(V):
    """transformer.py:398-405 at C3's real sizes: M = 16384 tokens, d = 1024, vocab padded to 50304."""
    from plaincv_b200 import ops
    torch.backends.cuda.matmul.allow_tf32 = False
    M, d, Vp = 16384, 1024, 50304
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn(M, d, device="cuda", generator=g).to(torch.bfloat16)
    w = torch.zeros(d, Vp, device="cuda", dtype=torch.bfloat16)
    w[:, :V] = (torch.randn(d, V, device="cuda", generator=g) * 0.02).to(torch.bfloat16)
    lib = _lib()
    p0 = lib.pcv_gemm_pair_launch_count()
    # forward: logits[M, Vp]
    y = ops.linear_fwd(x, w)
    ref = _fp32_matmul_chunks(x, w.float())
    e_fwd = rel_l2(y[:, :V], ref[:, :V])
    assert float(y[:, V:].float().abs().max()) == 0.0  # pad columns of the head stay exactly zero
    del ref
    # dgrad: dx[M, d] = dlogits[M, Vp] @ w^T, K = 50304 (dlogits is softmax-minus-onehot sized: ~1/M per entry)
    dl = torch.zeros(M, Vp, device="cuda", dtype=torch.bfloat16)
    dl[:, :V] = (torch.randn(M, V, device="cuda", generator=g) * (1.0 / M)).to(torch.bfloat16)
    dx = ops.linear_dgrad(dl, w)
    ref_dx = _fp32_matmul_chunks(dl, w.float().t().contiguous())
    e_dgrad = rel_l2(dx, ref_dx)
    # wgrad: dw[d, Vp] += x^T dlogits, split-K chosen by the library, fp32 reduce-add on top of existing contents
    dw0 = torch.randn(d, Vp, device="cuda", generator=g) * 1e-3
    dw = dw0.clone()
    ops.linear_wgrad(x, dl, dw)
    ref_dw = dw0 + x.float().t() @ dl.float()
    e_wgrad = rel_l2(dw, ref_dw)
    torch.cuda.synchronize()
    assert lib.pcv_gemm_pair_launch_count() - p0 == 3
    print(f"LM head V={V}: fwd rel-L2 {e_fwd:.2e}, dgrad {e_dgrad:.2e}, wgrad {e_wgrad:.2e}")
    assert e_fwd < 4e-3 and e_dgrad < 4e-3  # one bf16 rounding of an fp32-accumulated result: ~2^-9 / sqrt(3)
    assert e_wgrad < 1e-5


# ------------------------------------------------------------------------------------------------ csrc/test_gemm cases
_GEMM_CASES = [
    # (M, N, K, epilogue, splits)
    (128, 128, 64, 2, 1), (256, 256, 64, 2, 1), (128, 128, 256, 2, 1),
    (256, 384, 192, 0, 1),
    (200, 136, 72, 2, 1),       # ragged M / N / K tails
    (1000, 2736, 1000, 0, 1),
    (384, 520, 2000, 1, 4),     # split-K reduce-add into existing fp32 contents
    (3000, 5472, 520, 0, 1),    # several tiles per CTA / pair
    (4096, 1536, 256, 0, 1),    # 96 pair tiles: the last wave (22) runs as halved 256 x 128 tiles
]


@pytest.mark.parametrize("mode", [1, 2, 26], ids=["single_cta", "cta_pair", "cta_pair_dynamic"])
@pytest.mark.parametrize("a_mn,b_mn", [(0, 0), (0, 1), (1, 0), (1, 1)])
def test_gemm_cases_both_tilings(mode, a_mn, b_mn):
    """pcv_gemm_bf16 for every operand-major combination (forward = A K-major / B MN-major, dgrad = K / K, wgrad =
    MN / MN) under both tilings — the CTA-pair one with the static and with the dynamic (cluster-launch-control) tile
    schedule — against fp32 matmul of the same bf16 inputs.  ld = extent rounded up to 8 + 8 (TMA's 16-byte rule, and
    rows that do not end where the matrix does)."""
    from plaincv_b200 import ops
    torch.backends.cuda.matmul.allow_tf32 = False
    lib = _lib()
    g = torch.Generator(device="cuda").manual_seed(3)

    def pad8(n):
        return (n + 7) // 8 * 8 + 8

    dynamic = mode == 26
    if dynamic:
        mode = 2
        lib.pcv_gemm_set_mode(6)
    lib.pcv_gemm_set_mode(mode)
    try:
        for (M, N, K, epi, splits) in _GEMM_CASES:
            a_shape = (K, pad8(M)) if a_mn else (M, pad8(K))
            b_shape = (K, pad8(N)) if b_mn else (N, pad8(K))
            A = torch.randn(a_shape, device="cuda", generator=g).to(torch.bfloat16)
            Bm = torch.randn(b_shape, device="cuda", generator=g).to(torch.bfloat16)
            a_log = (A[:, :M].t() if a_mn else A[:, :K]).float()   # [M, K]
            b_log = (Bm[:, :N].t() if b_mn else Bm[:, :K]).float()  # [N, K]
            assert N % 8 == 0
            Np = N
            ldc = Np + 8
            alpha = 0.5
            if epi == 0:
                C = torch.full((M, ldc), 7.0, device="cuda", dtype=torch.bfloat16)
            else:
                C = torch.randn(M, ldc, device="cuda", generator=g)
            C0 = C.clone()
            p0, n0 = lib.pcv_gemm_pair_launch_count(), lib.pcv_launch_count()
            ops.gemm(A, Bm, C, M, Np, K, A.shape[1], Bm.shape[1], ldc, bool(a_mn), bool(b_mn), epi, splits, alpha)
            torch.cuda.synchronize()
            assert lib.pcv_launch_count() - n0 == 1
            assert lib.pcv_gemm_pair_launch_count() - p0 == (1 if mode == 2 else 0)
            want = alpha * (a_log @ b_log.t())
            if epi == 1:
                want = want + C0[:, :N]
            tol = 4e-3 if epi == 0 else 2e-5
            e = rel_l2(C[:, :N], want)
            assert e < tol, (M, N, K, a_mn, b_mn, epi, splits, mode, e)
            # nothing outside [M, Np) was written (columns N..Np are the zero-product of out-of-range B rows)
            assert torch.equal(C[:, Np:], C0[:, Np:]), (M, N, K, "wrote past the matrix")
    finally:
        lib.pcv_gemm_set_mode(0)
        lib.pcv_gemm_set_mode(6 if os.environ.get("PCV_GEMM_DYNAMIC", "0") not in ("", "0") else 5)
