"""SURVEY §8f N4 on the GPU: the tangent kernels of csrc/curvature.cu against fp32 forward-mode references
(torch.func.jvp of the same closed forms), and the GGN matvec J^T H J v of plaincv_b200.curvature against the
oracle's restatement of optim/ggn_utils.py:69-153 — plus the size-independent properties (symmetry, positive
semi-definiteness, linearity, null direction of the softmax Hessian) at the full C3 shapes.

Tolerances: single kernels rel. L2 <= 1e-2 against fp32 references on bf16-rounded inputs (like tests/test_kernels_gpu);
J v against the fp32 oracle <= 2e-2, G v per leaf <= max(3e-2, 1.25 x the oracle's bf16-emulating mode) and never
above 8e-2 (two bf16 passes — tangent and transpose — stack their rounding)."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import lm_oracle as O  # noqa: E402


def rel_l2(a, b):
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    return float((a - b).norm() / (b.norm() + 1e-30))


@pytest.fixture(scope="module")
def ops():
    from plaincv_b200 import ops as _ops
    return _ops


def _bf(x):
    return x.to(torch.bfloat16)


# ------------------------------------------------------------------------------------------------ single kernels
@pytest.mark.parametrize("M,d,with_delta", [(64, 256, True), (37, 1024, True), (16, 768, False), (8, 2048, True)])
def test_rmsnorm_residual_jvp(ops, M, d, with_delta):
    g = torch.Generator().manual_seed(M + d)
    res = _bf(torch.randn(M, d, generator=g))
    delta = _bf(torch.randn(M, d, generator=g)) if with_delta else None
    tres = _bf(torch.randn(M, d, generator=g))
    tdelta = _bf(torch.randn(M, d, generator=g)) if with_delta else None
    scale = 1.0 + 0.1 * torch.randn(d, generator=g)
    tscale = torch.randn(d, generator=g)
    eps = 1e-6
    res_out, y, rstd = ops.rmsnorm_residual_fwd(res.cuda(), None if delta is None else delta.cuda(), scale.cuda(), eps)
    tres_out, ty = ops.rmsnorm_residual_jvp(res_out, rstd, scale.cuda(), tscale.cuda(), tres.cuda(),
                                            None if tdelta is None else tdelta.cuda())
    x = (res.float() + delta.float()).to(torch.bfloat16).float() if with_delta else res.float()
    tx = (tres.float() + tdelta.float()).to(torch.bfloat16).float() if with_delta else tres.float()

    def f(xx, ss):
        return xx * torch.rsqrt((xx * xx).mean(-1, keepdim=True) + eps) * ss

    _, ty_ref = torch.func.jvp(f, (x, scale), (tx, tscale))
    assert rel_l2(ty, ty_ref) < 1e-2
    if with_delta:
        assert torch.equal(tres_out.cpu(), tx.to(torch.bfloat16))
    else:
        assert tres_out.data_ptr() == tres.cuda().data_ptr() or torch.equal(tres_out.cpu(), tres)


@pytest.mark.parametrize("M,hid", [(64, 704), (33, 2752)])
def test_swiglu_jvp(ops, M, hid):
    g = torch.Generator().manual_seed(hid)
    gu = _bf(torch.randn(M, 2 * hid, generator=g) * 2)
    tgu = _bf(torch.randn(M, 2 * hid, generator=g))
    th = ops.swiglu_jvp(gu.cuda(), tgu.cuda())

    def f(z):
        return torch.nn.functional.silu(z[:, :hid]) * z[:, hid:]

    _, ref = torch.func.jvp(f, (gu.float(),), (tgu.float(),))
    assert rel_l2(th, ref) < 1e-2


@pytest.mark.parametrize("mlp", ["mlp", "mlp_relu_sq"])
def test_act_jvp(ops, mlp):
    g = torch.Generator().manual_seed(3)
    x = _bf(torch.randn(40, 1408, generator=g) * 2)
    tx = _bf(torch.randn(40, 1408, generator=g))
    th = ops.act_jvp(x.cuda(), tx.cuda(), mlp)
    f = torch.nn.functional.silu if mlp == "mlp" else (lambda z: torch.square(torch.relu(z)))
    _, ref = torch.func.jvp(f, (x.float(),), (tx.float(),))
    assert rel_l2(th, ref) < 1e-2


@pytest.mark.parametrize("M,V,ldl", [(16, 1000, 1024), (5, 50257, 50304), (7, 50280, 50304),
                                     (3, 60001, 60032)])  # the last: a row too long for the shared-memory cache
def test_ce_hessian_vec(ops, M, V, ldl):
    """pcv_ce_hessian_vec against softmax_cross_entropy_hessian_vec (optim/ggn_utils.py:19-53) as restated by the
    oracle; pad columns must come out zero; in-place form identical."""
    g = torch.Generator().manual_seed(V)
    logits = torch.zeros(M, ldl)
    logits[:, :V] = torch.randn(M, V, generator=g) * 3
    vec = torch.zeros(M, ldl)
    vec[:, :V] = torch.randn(M, V, generator=g)
    lb, vb = _bf(logits), _bf(vec)
    scale = 1.0 / M
    hv = ops.ce_hessian_vec(lb.cuda(), vb.cuda(), V, scale)
    ref = O.softmax_cross_entropy_hessian_vec(lb[:, :V].float(), None, vb[:, :V].float()) * scale
    assert rel_l2(hv[:, :V], ref) < 1e-2
    assert float(hv[:, V:].float().abs().sum()) == 0.0
    # rows of H v sum to zero (1^T (diag(p) - p p^T) = 0) up to the bf16 rounding of the entries
    assert float(hv[:, :V].float().sum(-1).abs().max()) < 1e-2 * float(hv[:, :V].float().abs().sum(-1).max())
    v2 = vb.cuda().clone()
    out = ops.ce_hessian_vec(lb.cuda(), v2, V, scale, out=v2)
    assert out.data_ptr() == v2.data_ptr() and torch.equal(out, hv)
    # the public wrapper (un-padded last dimension that is a multiple of 8)
    if V % 8 == 0:
        from plaincv_b200.curvature import softmax_cross_entropy_hessian_vec
        hv2 = softmax_cross_entropy_hessian_vec(lb[:, :V].cuda(), None, vb[:, :V].cuda())
        assert rel_l2(hv2, ref / scale) < 1e-2


def _attn_ref(q, k, v, doc_start):
    """fp32 dot_product_attention on rotated q, k: [B, T, N, 64]; causal, optionally intersected with documents."""
    B, T, N, H = q.shape
    s = torch.einsum("btnh,bsnh->bnts", q, k) / math.sqrt(H)
    t = torch.arange(T)
    mask = (t[None, :] <= t[:, None])[None].expand(B, T, T)
    if doc_start is not None:
        mask = mask & (t[None, None, :] >= doc_start[:, :, None])
    s = torch.where(mask[:, None], s, torch.full((), -1e30))
    return torch.einsum("bnts,bsnh->btnh", torch.softmax(s, -1), v)


@pytest.mark.parametrize("B,T,NH,doc_mean", [(2, 128, 2, None), (1, 384, 3, None), (2, 512, 2, 100), (1, 1024, 1, 300),
                                             (1, 2048, 2, 512), (1, 4096, 1, None)])
def test_attention_jvp(ops, B, T, NH, doc_mean):
    """pcv_attn_jvp against torch.func.jvp of fp32 softmax attention on the same bf16-rounded q, k, v and tangents."""
    g = torch.Generator().manual_seed(T + NH)
    d = NH * 64
    qkv = _bf(torch.randn(B * T, 3 * d, generator=g))
    tqkv = _bf(torch.randn(B * T, 3 * d, generator=g))
    ds = None
    if doc_mean is not None:
        batch = O.synthetic_batch(B, T, 100, seed=T, doc_mean=doc_mean)
        from plaincv_b200.batch import prepare_batch
        _, _, ds = prepare_batch(batch, T, True)
        ds = ds.view(B, T)
    ds_dev = None if ds is None else ds.to(torch.int32).cuda().contiguous().view(-1)
    out, lse = ops.attn_fwd(qkv.cuda(), ds_dev, B, T, NH)
    tout = ops.attn_jvp(qkv.cuda(), tqkv.cuda(), ds_dev, out, lse, B, T, NH)

    def split(z):
        z = z.float().view(B, T, 3, NH, 64)
        return z[:, :, 0], z[:, :, 1], z[:, :, 2]

    q, k, v = split(qkv)
    tq, tk, tv = split(tqkv)
    o_ref, to_ref = torch.func.jvp(lambda a, b, c: _attn_ref(a, b, c, None if ds is None else ds.long()),
                                   (q, k, v), (tq, tk, tv))
    assert rel_l2(out.view(B, T, NH, 64), o_ref) < 1e-2
    assert rel_l2(tout.view(B, T, NH, 64), to_ref) < 1.5e-2


# ------------------------------------------------------------------------------------------------ whole path
def _build(cfg_o, seed=0):
    from plaincv_b200.model import ModelConfig, Transformer
    cfg = ModelConfig(vocab_size=cfg_o.vocab_size, seq_len=cfg_o.seq_len, dim=cfg_o.dim, expand=cfg_o.expand,
                      n_layers=cfg_o.n_layers, n_heads=cfg_o.n_heads, mlp=cfg_o.mlp, rope_theta=cfg_o.rope_theta,
                      tie_embeddings=cfg_o.tie_embeddings)
    model = Transformer(cfg, device="cuda", seed=seed)
    params = O.init_params(cfg_o, seed=seed)
    model.load_flax_params(params)
    return model, params


def _direction(params, seed):
    """A parameter-shaped direction with the scale of the parameters themselves (so every leaf contributes)."""
    g = torch.Generator().manual_seed(seed)
    return {k: torch.randn(v.shape, generator=g) * (0.02 if v.dim() > 1 else 0.1) for k, v in params.items()}


def _dot(a, b):
    return sum(float((a[k].double().cpu() * b[k].double().cpu()).sum()) for k in a)


@pytest.mark.parametrize("mlp,tied,use_doc,T", [("glu", False, False, 256), ("glu", False, True, 256), ("mlp", True, False, 256),
                                                ("mlp_relu_sq", False, False, 256), ("glu", False, True, 200)])
def test_jvp_and_ggn_matvec_against_oracle(mlp, tied, use_doc, T):
    """(the T = 200 case: a sequence length that is not a multiple of the attention tile runs right-padded)"""
    from plaincv_b200.curvature import TangentBuffers, jvp_logits, make_ggn_matvec_fn_lm
    V, B = 1000, 2
    cfg_o = O.ModelConfig(vocab_size=V, seq_len=T, dim=128, expand=8 / 3, n_layers=2, n_heads=2, mlp=mlp,
                          tie_embeddings=tied)
    model, params = _build(cfg_o, seed=5)
    batch = O.synthetic_batch(B, T, V, seed=21, doc_mean=70 if use_doc else None)
    inputs, labels, mask = O.prepare_batch(batch, T, use_doc)
    vec = _direction(params, 9)

    # J v
    tb = TangentBuffers(model)
    tb.load(vec)
    logits, tlogits = jvp_logits(model, inputs, tb, attn_mask=mask)
    lo_ref, tl_ref = torch.func.jvp(lambda p: O.forward(p, cfg_o, inputs, mask, torch.float32), (params,), (vec,))
    assert rel_l2(logits[:, :V], lo_ref.reshape(B * T, V)) < 1e-2
    e_jvp = rel_l2(tlogits[:, :V], tl_ref.reshape(B * T, V))
    assert float(tlogits[:, V:].float().abs().sum()) == 0.0
    assert e_jvp < 2e-2, e_jvp

    # G v = J^T H J v / (B*T)
    matvec = make_ggn_matvec_fn_lm(model, (inputs, labels, mask))
    gv = matvec(None, vec, None)
    torch.cuda.synchronize()
    ref = O.ggn_matvec(params, cfg_o, vec, inputs, labels, mask, torch.float32)
    emu = O.ggn_matvec(params, cfg_o, vec, inputs, labels, mask, torch.bfloat16)
    worst, bad = ("", 0.0, 0.0), []
    for k in ref:
        e, e_emu = rel_l2(gv[k], ref[k]), rel_l2(emu[k], ref[k])
        if e > worst[1]:
            worst = (k, e, e_emu)
        if e > max(3e-2, 1.25 * e_emu) or e > 8e-2:
            bad.append((k, e, e_emu))
    q, q_ref = _dot(vec, gv), _dot(vec, ref)
    print("J v rel-L2", e_jvp, "worst G v leaf (kernel, bf16-emulating oracle):", worst, "v^T G v", q, q_ref)
    assert not bad, bad
    assert q > 0 and abs(q - q_ref) / q_ref < 2e-2


def test_ggn_probe_batches_average():
    """input_ids [n_probe, B, T] -> mean of the per-probe products (optim/ggn_utils.py:137-150)."""
    from plaincv_b200.curvature import make_ggn_matvec_fn_lm
    V, T = 1000, 128
    cfg_o = O.ModelConfig(vocab_size=V, seq_len=T, dim=128, expand=8 / 3, n_layers=2, n_heads=2)
    model, params = _build(cfg_o, seed=6)
    ids = torch.randint(0, V, (2, 2, T), generator=torch.Generator().manual_seed(1))
    labels = torch.randint(0, V, (2, 2, T), generator=torch.Generator().manual_seed(2))
    vec = _direction(params, 3)
    both = make_ggn_matvec_fn_lm(model, (ids, labels, None))(None, vec)
    g0 = make_ggn_matvec_fn_lm(model, (ids[0], labels[0], None))(None, vec)
    g1 = make_ggn_matvec_fn_lm(model, (ids[1], labels[1], None))(None, vec)
    for k in both:
        assert rel_l2(both[k], 0.5 * (g0[k] + g1[k])) < 1e-2, k


def test_ggn_properties_full_size_c3_shapes():
    """C3 widths (d=1024, 16 heads, hid 2730, vocab 50280, seq 2048) with 2 layers and one sequence, where the oracle
    is too slow: through the public matvec G is symmetric (u^T G v = v^T G u), positive semi-definite and linear."""
    from plaincv_b200.curvature import make_ggn_matvec_fn_lm
    from plaincv_b200.model import ModelConfig, Transformer
    cfg = ModelConfig(vocab_size=50280, seq_len=2048, dim=1024, expand=8 / 3, n_layers=2, n_heads=16)
    model = Transformer(cfg, device="cuda", seed=0)
    params = model.export_flax_params()
    ids = torch.randint(0, 50280, (1, 2048), generator=torch.Generator().manual_seed(0))
    matvec = make_ggn_matvec_fn_lm(model, (ids, ids, None))
    u = {k: v.cpu() for k, v in _direction(params, 1).items()}
    v = {k: w.cpu() for k, w in _direction(params, 2).items()}
    gu, gv = matvec(None, u), matvec(None, v)
    uGv, vGu, vGv, uGu = _dot(u, gv), _dot(v, gu), _dot(v, gv), _dot(u, gu)
    print("u^T G v", uGv, "v^T G u", vGu, "v^T G v", vGv, "u^T G u", uGu)
    assert vGv > 0 and uGu > 0
    assert abs(uGv - vGu) <= 3e-2 * math.sqrt(vGv * uGu)
    g2v = matvec(None, {k: 2.0 * w for k, w in v.items()})
    for k in ("layers_0/attn/w_qkv/kernel", "layers_1/mlp/fc2/kernel", "lm_head/kernel", "embed_tokens/embedding"):
        assert rel_l2(g2v[k], 2.0 * gv[k]) < 2e-2, k


def test_ggn_matvec_against_golden_fixture():
    """The CUDA GGN matvec against tests/golden/small_ragged.json (oracle fp32 output committed by
    tools/make_golden.py: ragged vocab 1003, packed documents), without running the oracle: v^T G v within 2e-2,
    per-leaf norms within 3e-2, sampled entries within 8e-2 relative + 1.5e-1 of the leaf's rms."""
    import json
    import os
    from plaincv_b200.batch import prepare_batch, synthetic_batch
    from plaincv_b200.curvature import make_ggn_matvec_fn_lm
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "small_ragged.json")) as f:
        gold = json.load(f)
    spec = gold["spec"]
    cfg_o = O.ModelConfig(vocab_size=spec["vocab_size"], seq_len=spec["seq_len"], dim=spec["dim"], expand=8 / 3,
                          n_layers=spec["n_layers"], n_heads=spec["n_heads"])
    model, params = _build(cfg_o, seed=spec["seed"])
    batch = synthetic_batch(spec["batch"], spec["seq_len"], spec["vocab_size"], seed=spec["data_seed"],
                            doc_mean=spec["doc_mean"])
    inputs, labels, doc_start = prepare_batch(batch, spec["seq_len"], True)
    g = torch.Generator().manual_seed(gold["ggn"]["seed"])
    vec = {k: torch.randn(v.shape, generator=g) * (0.02 if v.dim() > 1 else 0.1) for k, v in params.items()}
    gv = make_ggn_matvec_fn_lm(model, (inputs, labels, doc_start))(None, vec)   # document starts instead of a mask
    vGv = _dot(vec, gv)
    assert abs(vGv - gold["ggn"]["vGv"]) <= 2e-2 * gold["ggn"]["vGv"], (vGv, gold["ggn"]["vGv"])
    for k, rec in gold["ggn"]["leaves"].items():
        got = gv[k].float().cpu()
        assert abs(float(got.double().norm()) - rec["norm"]) <= 3e-2 * rec["norm"], (k, float(got.norm()), rec["norm"])
        rms = rec["norm"] / math.sqrt(got.numel())
        flat = got.reshape(-1)
        for i, val in zip(rec["index"], rec["values"]):
            assert abs(float(flat[i]) - val) <= 8e-2 * abs(val) + 1.5e-1 * rms, (k, i, float(flat[i]), val)
