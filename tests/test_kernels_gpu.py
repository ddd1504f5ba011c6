"""Per-kernel parity: each CUDA entry point (called through the C ABI via plaincv_b200.ops) against the CPU oracle on
the same seeded inputs.  Tolerances: bf16 kernels vs the fp32 oracle evaluated on the same bf16-rounded inputs —
relative L2 <= 1e-2 (the north star's gradient tolerance), fp32 outputs tighter where stated."""
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import lm_oracle as O  # noqa: E402  (test infrastructure: the checker)


def rel_l2(a: torch.Tensor, b: torch.Tensor) -> float:
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    return float((a - b).norm() / (b.norm() + 1e-30))


@pytest.fixture(scope="module")
def ops():
    from plaincv_b200 import ops as _ops
    return _ops


def _bf(x):
    return x.to(torch.bfloat16)


@pytest.mark.parametrize("M,K,N", [(256, 256, 768), (512, 768, 2048), (384, 1024, 2736), (1024, 256, 1368)])
def test_linear_fwd_dgrad_wgrad(ops, M, K, N):
    g = torch.Generator().manual_seed(0)
    x = _bf(torch.randn(M, K, generator=g))
    w = _bf(torch.randn(K, N, generator=g) * 0.05)
    dy = _bf(torch.randn(M, N, generator=g))
    xd, wd, dyd = x.cuda(), w.cuda(), dy.cuda()
    y = ops.linear_fwd(xd, wd)
    assert rel_l2(y, x.float() @ w.float()) < 1e-2
    dx = ops.linear_dgrad(dyd, wd)
    assert rel_l2(dx, dy.float() @ w.float().t()) < 1e-2
    dw0 = torch.randn(K, N, generator=g)
    dwd = dw0.cuda()
    ops.linear_wgrad(xd, dyd, dwd)
    assert rel_l2(dwd, dw0 + x.float().t() @ dy.float()) < 1e-4  # fp32 accumulate-in-place (split-K red.add)


def test_cast_pad(ops):
    g = torch.Generator().manual_seed(1)
    src = torch.randn(37, 2730, generator=g)
    dst = torch.full((37, 2736), 7.0, dtype=torch.bfloat16).cuda()
    ops.cast_pad_bf16(src.cuda(), dst, 37, 2730, 2730, 2736)
    out = dst.cpu()
    assert torch.equal(out[:, :2730], src.to(torch.bfloat16))  # bit-exact round-to-nearest-even
    assert torch.equal(out[:, 2730:], torch.zeros(37, 6, dtype=torch.bfloat16))


def test_embed_fwd_bwd(ops):
    g = torch.Generator().manual_seed(2)
    V, d, M = 1000, 256, 777
    table = _bf(torch.randn(V, d, generator=g))
    ids = torch.randint(0, V, (M,), generator=g, dtype=torch.int32)
    ids[:50] = 3  # collisions: many rows scatter into the same table row
    x = ops.embed_fwd(ids.cuda(), table.cuda(), d)
    assert torch.equal(x.cpu(), table[ids.long()])  # a gather is bit-exact
    dx = _bf(torch.randn(M, d, generator=g))
    dE = torch.zeros(V, d).cuda()
    ops.embed_bwd(ids.cuda(), dx.cuda(), dE, d, d)
    want = torch.zeros(V, d).index_add_(0, ids.long(), dx.float())
    assert rel_l2(dE, want) < 1e-5


@pytest.mark.parametrize("M,d", [(2048, 256), (1024, 768), (4096, 1024)])
@pytest.mark.parametrize("with_delta", [False, True])
def test_rmsnorm_residual(ops, M, d, with_delta):
    g = torch.Generator().manual_seed(3)
    res = _bf(torch.randn(M, d, generator=g))
    delta = _bf(torch.randn(M, d, generator=g)) if with_delta else None
    scale = 1.0 + 0.1 * torch.randn(d, generator=g)
    eps = 1e-6
    res_out, y, rstd = ops.rmsnorm_residual_fwd(res.cuda(), None if delta is None else delta.cuda(), scale.cuda(), eps)
    r = res if delta is None else _bf(res.float() + delta.float())
    if with_delta:
        assert torch.equal(res_out.cpu(), r)  # residual stream: bf16(res + delta), bit-exact
    r32 = r.float().requires_grad_(True)
    sc = scale.clone().requires_grad_(True)
    y_ref = O.rmsnorm(r32, sc, eps, torch.float32)
    assert rel_l2(y, y_ref) < 4e-3
    ms = (r.float() ** 2).mean(-1)
    assert rel_l2(rstd, torch.rsqrt(ms + eps)) < 1e-5
    # backward
    dy = _bf(torch.randn(M, d, generator=g))
    dres_out = _bf(torch.randn(M, d, generator=g))
    y_ref.backward(dy.float())
    dscale = torch.zeros(d).cuda()
    dres_in = ops.rmsnorm_residual_bwd(dy.cuda(), dres_out.cuda(), res_out, scale.cuda(), rstd, dscale)
    assert rel_l2(dres_in, r32.grad + dres_out.float()) < 5e-3
    assert rel_l2(dscale, sc.grad) < 1e-3


@pytest.mark.parametrize("M,hid", [(512, 2736), (1024, 2048), (256, 688)])
def test_swiglu(ops, M, hid):
    g = torch.Generator().manual_seed(4)
    gu = _bf(torch.randn(M, 2 * hid, generator=g) * 2)
    dh = _bf(torch.randn(M, hid, generator=g))
    h = ops.swiglu_fwd(gu.cuda())
    gu32 = gu.float().requires_grad_(True)
    ref = O.silu(gu32[:, :hid]) * gu32[:, hid:]
    assert rel_l2(h, ref) < 4e-3
    ref.backward(dh.float())
    dgu = ops.swiglu_bwd(gu.cuda(), dh.cuda())
    assert rel_l2(dgu, gu32.grad) < 5e-3


@pytest.mark.parametrize("M,d,hid", [(512, 256, 688), (768, 1024, 2736), (300, 256, 128), (4096, 512, 1368)])
def test_fused_swiglu_gemms(ops, M, d, hid):
    """pcv_gemm_bf16_swiglu / pcv_gemm_bf16_dswiglu (GLU, transformer.py:110-134): against the oracle's fp32
    silu(x Wg) * (x Wu) and its autograd transpose, and bit-identical to the un-fused GEMM + SwiGLU kernels."""
    g = torch.Generator().manual_seed(21)
    x = _bf(torch.randn(M, d, generator=g))
    w_gu = _bf(torch.randn(d, 2 * hid, generator=g) * 0.05)
    w_down = _bf(torch.randn(hid, d, generator=g) * 0.05)
    dy = _bf(torch.randn(M, d, generator=g))
    xd, wgud, wdd, dyd = x.cuda(), w_gu.cuda(), w_down.cuda(), dy.cuda()
    gu, h = ops.linear_fwd_swiglu(xd, wgud)
    gu_ref = ops.linear_fwd(xd, wgud)
    assert torch.equal(gu, gu_ref)
    assert torch.equal(h, ops.swiglu_fwd(gu_ref))
    gu32 = (x.float() @ w_gu.float())
    assert rel_l2(gu, gu32) < 5e-3
    assert rel_l2(h, O.silu(gu32[:, :hid]) * gu32[:, hid:]) < 8e-3
    dgu = ops.linear_dgrad_dswiglu(dyd, wdd, gu)
    dgu_ref = ops.swiglu_bwd(gu, ops.linear_dgrad(dyd, wdd))
    assert torch.equal(dgu, dgu_ref)
    gul = gu.float().cpu().requires_grad_(True)
    hl = O.silu(gul[:, :hid]) * gul[:, hid:]
    hl.backward(dy.float() @ w_down.float().t())
    assert rel_l2(dgu, gul.grad) < 8e-3


def _attention_oracle(qkv, B, T, NH, theta, doc_start=None):
    """RoPE + attention on bf16-rounded qkv with the oracle's fp32 functions (autograd-enabled)."""
    d = NH * 64
    qkv3 = qkv.view(B, T, 3 * d)
    q, k, v = torch.split(qkv3, d, dim=-1)
    q = q.reshape(B, T, NH, 64)
    k = k.reshape(B, T, NH, 64)
    v = v.reshape(B, T, NH, 64)
    freqs = O.precompute_freqs_cis(64, T, theta)
    q, k = O.apply_rotary_emb_complex_like(q, k, freqs)
    if doc_start is None:
        out = O.dot_product_attention(q, k, v, None, True)
    else:
        s = torch.arange(T)
        mask = (s[None, None, :] >= doc_start[:, :, None]) & (s[None, None, :] <= s[None, :, None])
        out = O.dot_product_attention(q, k, v, mask[:, None], False)
    return out.reshape(B * T, d)


def _doc_starts(B, T, seed, mean):
    batch = O.synthetic_batch(B, T, 100, seed=seed, doc_mean=mean)
    from plaincv_b200.batch import doc_lengths_to_doc_start
    return torch.from_numpy(doc_lengths_to_doc_start(batch["docs_lengths"], T))


@pytest.mark.parametrize("B,T,NH,doc_mean", [(2, 256, 4, None), (1, 1024, 2, None), (2, 512, 3, 150), (1, 1024, 2, 40),
                                             (1, 128, 1, None), (2, 384, 2, 1000),
                                             # BASELINE's full sequence lengths: C3 / C5 (2048, packed documents of
                                             # mean 512) and C4 (4096), one or two heads so the fp32 oracle stays cheap
                                             (1, 2048, 2, None), (1, 2048, 2, 512), (1, 4096, 1, None), (1, 4096, 1, 700)])
def test_attention_fwd_bwd(ops, B, T, NH, doc_mean):
    g = torch.Generator().manual_seed(5)
    d = NH * 64
    theta = 500000.0
    qkv = _bf(torch.randn(B * T, 3 * d, generator=g))
    dout = _bf(torch.randn(B * T, d, generator=g))
    doc_start = _doc_starts(B, T, 11, doc_mean) if doc_mean else None
    cos_sin = ops.rope_table(64, T, theta, "cuda")
    ds_dev = None if doc_start is None else doc_start.to(torch.int32).cuda().view(-1)
    qkv_rot = ops.rope_qk_(qkv.cuda().clone(), cos_sin, B, T, NH)  # stand-alone form of the fused GEMM epilogue
    out, lse = ops.attn_fwd(qkv_rot, ds_dev, B, T, NH)
    torch.cuda.synchronize()
    qkv32 = qkv.float().requires_grad_(True)
    ref = _attention_oracle(qkv32, B, T, NH, theta, doc_start)
    assert rel_l2(out, ref) < 1e-2, "attention forward"
    ref.backward(dout.float())
    dqkv = ops.attn_bwd(qkv_rot, cos_sin, ds_dev, out, dout.cuda(), lse, B, T, NH)
    torch.cuda.synchronize()
    gref = qkv32.grad
    for name, sl in (("dq", slice(0, d)), ("dk", slice(d, 2 * d)), ("dv", slice(2 * d, 3 * d))):
        assert rel_l2(dqkv[:, sl], gref[:, sl]) < 1.5e-2, f"attention backward {name}"


def test_fused_rope_gemm_matches_gemm_then_rope(ops):
    """pcv_gemm_bf16_rope == RoPE(fp32 x@w) on the q|k columns, plain GEMM on the v columns (embedding.py:28-67)."""
    g = torch.Generator().manual_seed(12)
    B, T, NH = 2, 256, 4
    d = NH * 64
    x = _bf(torch.randn(B * T, d, generator=g))
    w = _bf(torch.randn(d, 3 * d, generator=g) * 0.05)
    cos_sin = ops.rope_table(64, T, 500000.0, "cuda")
    y = ops.linear_fwd_rope(x.cuda(), w.cuda(), cos_sin, T, 2 * d)
    y32 = (x.float() @ w.float()).view(B, T, 3 * d)
    q, k, v = torch.split(y32, d, dim=-1)
    freqs = O.precompute_freqs_cis(64, T, 500000.0)
    qr, kr = O.apply_rotary_emb_complex_like(q.reshape(B, T, NH, 64), k.reshape(B, T, NH, 64), freqs)
    ref = torch.cat([qr.reshape(B, T, d), kr.reshape(B, T, d), v], dim=-1).view(B * T, 3 * d)
    assert rel_l2(y, ref) < 5e-3
    # and the stand-alone rotation agrees with the fused one up to the extra bf16 rounding
    y2 = ops.rope_qk_(ops.linear_fwd(x.cuda(), w.cuda()), cos_sin, B, T, NH)
    assert rel_l2(y2, ref) < 8e-3


def test_attention_first_row_is_v0(ops):
    """Closed form: causal row 0 attends only to itself -> out[0] = v[0] (and doc starts behave the same)."""
    g = torch.Generator().manual_seed(6)
    B, T, NH = 1, 128, 2
    d = NH * 64
    qkv = _bf(torch.randn(B * T, 3 * d, generator=g))
    cos_sin = ops.rope_table(64, T, 500000.0, "cuda")
    ds = torch.arange(T, dtype=torch.int32).cuda()  # every token its own document
    out, _ = ops.attn_fwd(ops.rope_qk_(qkv.cuda().clone(), cos_sin, B, T, NH), ds, B, T, NH)
    assert torch.equal(out.cpu(), qkv[:, 2 * d:])  # each row returns exactly its own v


@pytest.mark.parametrize("M,V", [(256, 50257), (512, 50280), (128, 1000), (64, 1003)])
def test_softmax_ce(ops, M, V):
    g = torch.Generator().manual_seed(7)
    ldl = (V + 63) // 64 * 64  # the layout's row stride (model._pad64); pad columns hold garbage the kernel must ignore
    logits = _bf(torch.randn(M, ldl, generator=g) * 3)
    logits[:, V:] = 1e4
    labels = torch.randint(0, V, (M,), generator=g, dtype=torch.int32)
    logits[5, labels[5]] = 40.0  # a confidently-correct row
    l32 = logits[:, :V].float().requires_grad_(True)
    loss_ref, acc_ref = O.loss_and_acc(l32, labels)
    loss_ref.backward()
    buf = logits.cuda().clone()
    out2, dlog = ops.softmax_ce(buf, labels.cuda(), V, 1.0 / M, True)
    out2 = out2.cpu()
    assert abs(float(out2[0]) - float(loss_ref)) / float(loss_ref) < 1e-5
    assert float(out2[1]) == pytest.approx(float(acc_ref), abs=1e-7)
    assert rel_l2(dlog[:, :V], l32.grad) < 4e-3
    if ldl > V:
        assert float(dlog[:, V:].abs().max()) == 0.0
    # uniform logits: loss = ln V (closed form)
    buf = torch.zeros(M, ldl, dtype=torch.bfloat16).cuda()
    out2, _ = ops.softmax_ce(buf, labels.cuda(), V, 1.0, False)
    assert float(out2[0]) == pytest.approx(math.log(V), rel=1e-6)


@pytest.mark.parametrize("n,max_norm", [(1 << 20, 1.0), (1000003, 0.05), (4099, None), (3 * 1024 * 1024 + 8, 100.0)])
def test_clip_adamw_flat(ops, n, max_norm):
    """pcv_sumsq_partials + pcv_adamw_step against the oracle's _clip_grads + optax.adamw restatement
    (train_lm.py:173-178, 316-353; optim/factory.py:193-205) over three consecutive updates of one flat buffer;
    fp32 state within 2e-6 relative (same formula, different association), the norm within 1e-6, and the bf16 compute
    copy bit-identical to the rounding of the kernel's own fp32 parameters."""
    g = torch.Generator().manual_seed(5)
    p0 = torch.randn(n, generator=g) * 0.02
    lr, b1, b2, eps, wd = 3e-3, 0.9, 0.999, 1e-8, 0.1
    p_d, m_d, v_d = p0.cuda(), torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    pb = torch.zeros(n, dtype=torch.bfloat16, device="cuda")
    partials = torch.zeros(296, device="cuda")
    norm_d = torch.zeros(1, device="cuda")
    p_o = {"w": p0.clone()}
    m_o, v_o = {"w": torch.zeros(n)}, {"w": torch.zeros(n)}
    for step in (1, 2, 3):
        gr = torch.randn(n, generator=g) * (0.5 if step == 2 else 0.003)
        ops.sumsq_partials(gr.cuda(), partials)
        ops.adamw_step(p_d, gr.cuda(), m_d, v_d, pb, partials if max_norm is not None else None, norm_d,
                       max_norm if max_norm is not None else 0.0, lr, b1, b2, eps, wd, step)
        gc = O.clip_grads({"w": gr}, max_norm)
        p_o, m_o, v_o = O.adamw_step(p_o, gc, m_o, v_o, step, lr, b1, b2, eps, wd)
        torch.cuda.synchronize()
        if max_norm is not None:
            assert abs(float(norm_d) - float(gr.double().norm())) <= 1e-6 * float(gr.double().norm())
        assert rel_l2(p_d, p_o["w"]) < 2e-6
        assert rel_l2(m_d, m_o["w"]) < 2e-6
        assert rel_l2(v_d, v_o["w"]) < 2e-6
        assert float((p_d.cpu() - p_o["w"]).abs().max()) < 1e-6
        assert torch.equal(pb, p_d.to(torch.bfloat16))


@pytest.mark.parametrize("mlp", ["mlp", "mlp_relu_sq"])
@pytest.mark.parametrize("M,hid", [(512, 2736), (300, 688)])
def test_ungated_activations(ops, mlp, M, hid):
    """pcv_act_fwd / pcv_act_bwd: silu (MLP, transformer.py:70-97) and relu^2 (MLPReluSquared,
    :138-165) of a bf16 Dense output against the oracle's fp32 functions and their autograd transposes."""
    g = torch.Generator().manual_seed(31)
    x = _bf(torch.randn(M, hid, generator=g) * 1.5)
    dh = _bf(torch.randn(M, hid, generator=g))
    h = ops.act_fwd(x.cuda(), mlp)
    dx = ops.act_bwd(x.cuda(), dh.cuda(), mlp)
    xl = x.float().requires_grad_(True)
    ref = O.silu(xl) if mlp == "mlp" else torch.square(torch.relu(xl))
    ref.backward(dh.float())
    assert rel_l2(h, ref) < 4e-3
    assert rel_l2(dx, xl.grad) < 4e-3
    if mlp == "mlp_relu_sq":  # exact where the input is non-positive
        assert float(h.cpu()[x <= 0].abs().max()) == 0.0 and float(dx.cpu()[x <= 0].abs().max()) == 0.0


@pytest.mark.parametrize("B,T,NH", [(2, 1024, 4), (1, 2048, 16)])
def test_attention_against_flash_attn_library(ops, B, T, NH):
    """Independent GPU checker: pcv_attn_fwd / pcv_attn_bwd against flash_attn_func — a causal
    attention neither this repository nor its oracle wrote — on the same already-rotated bf16 q, k, v.  flash_attn is
    test infrastructure here (never imported by plaincv_b200); both sides are bf16 kernels, so they are compared with
    each other at 1e-2 and each sits within 1e-2 of the fp32 oracle in test_attention_fwd_bwd."""
    flash_attn = pytest.importorskip("flash_attn")
    g = torch.Generator().manual_seed(9)
    d = NH * 64
    qkv = _bf(torch.randn(B * T, 3 * d, generator=g)).cuda()
    dout = _bf(torch.randn(B * T, d, generator=g)).cuda()
    # identity rotation table (cos 1, sin 0): the un-rotation inside pcv_attn_bwd is then the identity as well, so the
    # comparison isolates the attention arithmetic (RoPE has its own tests above)
    cos_sin = torch.zeros(T, 32, 2, device="cuda")
    cos_sin[..., 0] = 1.0
    out, lse = ops.attn_fwd(qkv, None, B, T, NH)
    dqkv = ops.attn_bwd(qkv, cos_sin, None, out, dout, lse, B, T, NH)
    q, k, v = (t.reshape(B, T, NH, 64).detach().clone().requires_grad_(True) for t in qkv.view(B, T, 3 * d).split(d, -1))
    try:
        ref = flash_attn.flash_attn_func(q, k, v, causal=True)
        ref.backward(dout.view(B, T, NH, 64))
    except RuntimeError as e:  # a flash-attn build without kernels for this GPU is not this repository's failure
        pytest.skip(f"flash_attn cannot run here: {e}")
    assert rel_l2(out, ref.reshape(B * T, d)) < 1e-2
    for name, sl, gr in (("dq", slice(0, d), q.grad), ("dk", slice(d, 2 * d), k.grad), ("dv", slice(2 * d, 3 * d), v.grad)):
        assert rel_l2(dqkv[:, sl], gr.reshape(B * T, d)) < 1.5e-2, name


@pytest.mark.parametrize("M,V,tied", [(256, 50257, False), (512, 50280, False), (128, 1000, False), (192, 1003, True),
                                      (16384, 50280, False)])
def test_head_fused_with_cross_entropy(ops, M, V, tied):
    """LM head fused with softmax cross-entropy (transformer.py:398-405 + train_lm.py:181-186): pcv_gemm_bf16_ce leaves
    the per-(row, 128-column slot) online-softmax partials in the head GEMM's epilogue, pcv_ce_from_stats finishes in one
    pass.  Against (a) the un-fused pair pcv_gemm_bf16 + pcv_softmax_ce — identical logits, loss / accuracy / d(logits)
    equal up to fp32 summation order — and (b) the oracle's loss_and_acc on the same bf16 logits."""
    class _Head:
        pass
    g = torch.Generator(device="cuda").manual_seed(21)
    d = 256
    vp = (V + 63) // 64 * 64
    hf = torch.randn(M, d, device="cuda", generator=g).to(torch.bfloat16)
    head = _Head()
    head.tied, head.vocab, head.v_pad = tied, V, vp
    if tied:  # table [V_pad, d], pad rows zero
        w = torch.zeros(vp, d, device="cuda", dtype=torch.bfloat16)
        w[:V] = (torch.randn(V, d, device="cuda", generator=g) * 0.3).to(torch.bfloat16)
    else:     # lm_head kernel [d, V_pad], pad columns zero
        w = torch.zeros(d, vp, device="cuda", dtype=torch.bfloat16)
        w[:, :V] = (torch.randn(d, V, device="cuda", generator=g) * 0.3).to(torch.bfloat16)
    head.w_lm = w
    labels = torch.randint(0, V, (M,), device="cuda", generator=g, dtype=torch.int32)
    labels[0] = V - 1
    scale = 1.0 / M
    ref_logits = ops.head_logits(hf, head)
    logits, ce_ms = ops.head_logits_ce(hf, head)
    assert torch.equal(logits, ref_logits)  # the fused epilogue stores the same bf16 logits
    lf = ref_logits[:, :V].float()
    out2_ref, dl_ref = ops.softmax_ce(ref_logits.clone(), labels, V, scale, True)
    out2, dl = ops.ce_from_stats(logits, labels, ce_ms, V, scale)
    torch.cuda.synchronize()
    assert abs(float(out2[0]) - float(out2_ref[0])) < 2e-6 * abs(float(out2_ref[0]))
    assert float(out2[1]) == float(out2_ref[1])
    assert rel_l2(dl, dl_ref) < 2e-3  # bf16 outputs of two fp32 computations that differ in summation order
    if vp > V:
        assert float(dl[:, V:].float().abs().max()) == 0.0
    if M <= 512:  # and against the oracle (fp32 softmax cross-entropy of the same bf16 logits, first-max argmax)
        lo = lf.cpu().requires_grad_(True)
        loss_o, acc_o = O.loss_and_acc(lo[None], labels.cpu()[None])
        loss_o.backward()
        assert abs(float(out2[0]) - float(loss_o)) < 1e-5 * float(loss_o)
        assert float(out2[1]) == pytest.approx(float(acc_o), abs=1e-7)
        assert rel_l2(dl[:, :V], lo.grad) < 6e-3


def test_model_with_fused_head_ce_matches_unfused(ops, monkeypatch):
    """The opt-in model path (PCV_HEAD_CE_FUSED=1): same loss, gradients within the run-to-run noise of the default."""
    from plaincv_b200.engine import Engine
    from plaincv_b200.model import ModelConfig, Transformer
    cfg = ModelConfig(vocab_size=50257, seq_len=256, dim=256, expand=8 / 3, n_layers=2, n_heads=4)
    model = Transformer(cfg, device="cuda", seed=0)
    batch = O.synthetic_batch(4, 256, 50257, seed=3)
    eng = Engine(model, seq_len=256)
    monkeypatch.setattr(ops, "_HEAD_CE_FUSED", False)
    l0, a0 = eng.compute_grads(batch)
    g0 = model.flat_grad.clone()
    monkeypatch.setattr(ops, "_HEAD_CE_FUSED", True)
    n0 = _lib_count()
    l1, a1 = eng.compute_grads(batch)
    assert abs(float(l1) - float(l0)) < 2e-6 * float(l0) and float(a1) == float(a0)
    assert rel_l2(model.flat_grad, g0) < 2e-3


def _lib_count():
    from plaincv_b200 import _lib
    return _lib.load().pcv_launch_count()
