"""CPU suite, part 1: the oracle (oracle/lm_oracle.py) against closed forms, against an independent numpy
re-derivation, against fp64 finite differences and against the committed golden fixtures (tests/golden/*.json, made
by tools/make_golden.py).  The reference has no golden vectors of its own: parity with it is
"unpinned"; what is pinned here is that the restatement does what the cited reference lines say."""
import json
import math
import os

import numpy as np
import pytest
import torch

from oracle import lm_oracle as O

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


# ------------------------------------------------------------------------------------------------ RoPE (embedding.py)
def test_freqs_cis_shape_and_values():
    f = O.precompute_freqs_cis(64, 16, theta=500000.0)
    assert tuple(f.shape) == (1, 16, 1, 32, 2)
    # position 0: angle 0 -> cos 1, sin 0
    assert torch.all(f[0, 0, 0, :, 0] == 1) and torch.all(f[0, 0, 0, :, 1] == 0)
    # pair i at position t: angle t * theta^(-2i/dim)
    t, i = 5, 3
    ang = t * 500000.0 ** (-2 * i / 64)
    assert abs(float(f[0, t, 0, i, 0]) - math.cos(ang)) < 1e-6
    assert abs(float(f[0, t, 0, i, 1]) - math.sin(ang)) < 1e-6
    with pytest.raises(ValueError):
        O.precompute_freqs_cis(63, 4)


def test_rope_is_interleaved_pair_rotation():
    g = torch.Generator().manual_seed(0)
    q = torch.randn(2, 8, 3, 64, generator=g)
    k = torch.randn(2, 8, 3, 64, generator=g)
    f = O.precompute_freqs_cis(64, 8, theta=10000.0)
    qr, kr = O.apply_rotary_emb_complex_like(q, k, f)
    # independent derivation with complex numbers: (x[2i] + i x[2i+1]) * exp(i * angle)
    ang = np.outer(np.arange(8), 10000.0 ** (-np.arange(0, 64, 2) / 64))
    rot = np.exp(1j * ang)[None, :, None, :]
    for x, xr in ((q, qr), (k, kr)):
        z = x.numpy()[..., 0::2] + 1j * x.numpy()[..., 1::2]
        zr = z * rot
        exp = np.stack([zr.real, zr.imag], axis=-1).reshape(x.shape)
        np.testing.assert_allclose(xr.numpy(), exp, rtol=1e-5, atol=1e-5)
    # rotations preserve the norm of every pair; position 0 is the identity
    assert torch.allclose(qr.norm(dim=-1), q.norm(dim=-1), rtol=1e-5)
    assert torch.allclose(qr[:, 0], q[:, 0])
    # relative-position property: <rope(q,t), rope(k,s)> depends on t - s only
    qq = q[:, :1].expand(-1, 8, -1, -1).contiguous()
    kk = k[:, :1].expand(-1, 8, -1, -1).contiguous()
    a, b = O.apply_rotary_emb_complex_like(qq, kk, f)
    d1 = (a[:, 3] * b[:, 1]).sum(-1)
    d2 = (a[:, 6] * b[:, 4]).sum(-1)
    assert torch.allclose(d1, d2, atol=1e-4)


# ------------------------------------------------------------------------------------------------ layers
def test_rmsnorm_closed_form():
    x = torch.tensor([[3.0, -4.0, 0.0, 0.0]])
    scale = torch.tensor([1.0, 2.0, 3.0, 4.0])
    y = O.rmsnorm(x, scale, 1e-6, torch.float32)
    r = 1.0 / math.sqrt(25.0 / 4 + 1e-6)
    np.testing.assert_allclose(y.numpy(), [[3 * r, -8 * r, 0, 0]], rtol=1e-6)
    # no mean-centring: a constant row maps to +-scale
    y = O.rmsnorm(torch.full((1, 8), 5.0), torch.ones(8), 0.0, torch.float32)
    assert torch.allclose(y, torch.ones(1, 8))


def test_dense_is_in_out_layout_and_bf16_mode_rounds():
    x = torch.tensor([[1.0, 2.0]])
    w = torch.tensor([[1.0, 0.0, 2.0], [0.0, 1.0, 3.0]])  # (in=2, out=3)
    assert torch.equal(O.dense(x, w, torch.float32), torch.tensor([[1.0, 2.0, 8.0]]))
    x = torch.tensor([[1.0 + 2 ** -9]])  # not representable in bf16
    y = O.dense(x, torch.tensor([[1.0]]), torch.bfloat16)
    assert y.dtype == torch.bfloat16 and float(y) == 1.0


def test_attention_causal_closed_forms():
    g = torch.Generator().manual_seed(1)
    B, T, N, H = 2, 6, 2, 8
    q = torch.randn(B, T, N, H, generator=g)
    k = torch.randn(B, T, N, H, generator=g)
    v = torch.randn(B, T, N, H, generator=g)
    out = O.dot_product_attention(q, k, v, None, True)
    # row 0 sees only key 0
    assert torch.allclose(out[:, 0], v[:, 0], atol=1e-6)
    # zero queries -> uniform average over the visible prefix
    out0 = O.dot_product_attention(torch.zeros_like(q), k, v, None, True)
    for t in range(T):
        assert torch.allclose(out0[:, t], v[:, : t + 1].mean(1), atol=1e-6)
    # explicit softmax in numpy with scale 1/sqrt(H)
    s = np.einsum("btnh,bsnh->bnts", q.numpy(), k.numpy()) / math.sqrt(H)
    s = np.where(np.tril(np.ones((T, T), bool))[None, None], s, -np.inf)
    p = np.exp(s - s.max(-1, keepdims=True))
    p /= p.sum(-1, keepdims=True)
    exp = np.einsum("bnts,bsnh->btnh", p, v.numpy())
    np.testing.assert_allclose(out.numpy(), exp, rtol=1e-5, atol=1e-5)


def test_bool_mask_equals_causal_when_single_document():
    g = torch.Generator().manual_seed(2)
    q, k, v = (torch.randn(1, 5, 1, 4, generator=g) for _ in range(3))
    m = O.intra_doc_causal_mask([5], 5)[None, None]
    a = O.dot_product_attention(q, k, v, m, False)
    b = O.dot_product_attention(q, k, v, None, True)
    assert torch.allclose(a, b)


def test_intra_doc_mask_and_trim():
    m = O.intra_doc_causal_mask([2, 3], 5)
    exp = torch.tensor([[1, 0, 0, 0, 0], [1, 1, 0, 0, 0], [0, 0, 1, 0, 0], [0, 0, 1, 1, 0], [0, 0, 1, 1, 1]],
                       dtype=torch.bool)
    assert torch.equal(m, exp)
    with pytest.raises(ValueError):
        O.intra_doc_causal_mask([2, 2], 5)
    assert O.trim_last_token([3, 2]) == [3, 1]
    assert O.trim_last_token([3, 1]) == [3]
    assert O.trim_last_token([]) == []
    batch = {"input_ids": np.zeros((1, 6), np.int32), "docs_lengths": [[2, 4]]}
    mask = O.build_attn_mask(batch, 5)
    assert torch.equal(mask[0], O.intra_doc_causal_mask([2, 3], 5))
    with pytest.raises(ValueError):
        O.build_attn_mask({"input_ids": np.zeros((1, 6), np.int32)}, 5)
    with pytest.raises(ValueError):
        O.build_attn_mask({"input_ids": np.zeros((1, 7), np.int32), "docs_lengths": [[2, 5]]}, 5)
    with pytest.raises(ValueError):
        O.build_attn_mask({"input_ids": np.zeros((1, 6), np.int32), "docs_lengths": [[2, 5]]}, 5)


def test_loss_and_acc_closed_forms():
    V = 11
    logits = torch.zeros(2, 3, V)
    labels = torch.zeros(2, 3, dtype=torch.int32)
    loss, acc = O.loss_and_acc(logits, labels)
    assert abs(float(loss) - math.log(V)) < 1e-6
    assert float(acc) == 1.0  # ties: argmax returns the first maximum (index 0)
    labels = torch.full((2, 3), 4, dtype=torch.int32)
    assert float(O.loss_and_acc(logits, labels)[1]) == 0.0
    logits = torch.zeros(1, 1, V)
    logits[0, 0, 4] = 50.0
    loss, acc = O.loss_and_acc(logits, torch.tensor([[4]]))
    assert float(loss) < 1e-6 and float(acc) == 1.0


def test_param_tree_names_shapes_and_count():
    cfg = O.ModelConfig(vocab_size=50280, seq_len=2048, dim=1024, expand=float(8 / 3), n_layers=24, n_heads=16)
    assert cfg.hidden_dim == 2730  # int(float(Fraction("8/3")) * 1024), constructor.py:87 / transformer.py:296
    shapes = O.param_shapes(cfg)
    assert shapes["layers_3/attn/w_qkv/kernel"] == (1024, 3072)
    assert shapes["layers_3/mlp/fc2/kernel"] == (2730, 1024)
    assert shapes["lm_head/kernel"] == (1024, 50280)
    n = sum(int(np.prod(s)) for s in shapes.values())
    assert n == 404_964_352
    tiny = O.ModelConfig(vocab_size=50257, seq_len=256, dim=256, expand=8 / 3, n_layers=4, n_heads=4)
    assert tiny.hidden_dim == 682
    assert sum(int(np.prod(s)) for s in O.param_shapes(tiny).values()) == 28_877_568


def test_model_config_from_yaml_keys():
    cfg = O.model_config_from_cfg(dict(model="transformer", vocab_size=50280, d_model=1024, expand="8/3", n_layers=24,
                                       n_heads=16, mlp_class="glu", seq_len=2048))
    assert (cfg.dim, cfg.hidden_dim, cfg.rope_theta, cfg.rmsnorm_eps) == (1024, 2730, 500000.0, 1e-6)
    with pytest.raises(NotImplementedError):
        O.model_config_from_cfg(dict(model="pythia"))


# ------------------------------------------------------------------------------------------------ gradients
def _tiny(seed=0, vocab=37, mlp="glu"):
    cfg = O.ModelConfig(vocab_size=vocab, seq_len=12, dim=16, expand=8 / 3, n_layers=2, n_heads=2, mlp=mlp)
    params = O.init_params(cfg, seed=seed)
    params = {k: (v * 5).double() for k, v in params.items()}  # larger weights: gradients well above fd noise
    return cfg, params


@pytest.mark.parametrize("use_doc", [False, True])
def test_fp64_finite_difference_gradcheck(use_doc):
    cfg, params = _tiny()
    batch = O.synthetic_batch(2, 12, cfg.vocab_size, seed=3, doc_mean=4 if use_doc else None)
    inputs, labels, mask = O.prepare_batch(batch, 12, use_doc)
    grads, loss, _ = O.compute_grads(params, cfg, inputs, labels, mask, torch.float64)
    rng = np.random.default_rng(0)
    eps = 1e-6
    for name in ["embed_tokens/embedding", "layers_0/attn/w_qkv/kernel", "layers_1/attn/w_out/kernel",
                 "layers_0/mlp/fc_gate/kernel", "layers_1/mlp/fc2/kernel", "layers_1/mlp_norm/RMSNorm_0/scale",
                 "out_norm/RMSNorm_0/scale", "lm_head/kernel"]:
        p = params[name]
        for _ in range(3):
            idx = tuple(int(rng.integers(0, s)) for s in p.shape)
            if name == "embed_tokens/embedding":
                idx = (int(inputs[0, 0]), idx[1])  # a row that is actually gathered
            old = float(p[idx])
            p[idx] = old + eps
            lp = float(O.loss_and_acc(O.forward(params, cfg, inputs, mask, torch.float64), labels)[0])
            p[idx] = old - eps
            lm = float(O.loss_and_acc(O.forward(params, cfg, inputs, mask, torch.float64), labels)[0])
            p[idx] = old
            fd = (lp - lm) / (2 * eps)
            an = float(grads[name][idx])
            assert abs(fd - an) <= 1e-6 + 1e-4 * abs(an), (name, idx, fd, an)


def test_bf16_mode_is_close_to_fp32_mode():
    cfg = O.ModelConfig(vocab_size=257, seq_len=64, dim=128, expand=8 / 3, n_layers=2, n_heads=2)
    params = O.init_params(cfg, seed=0)
    batch = O.synthetic_batch(2, 64, 257, seed=9)
    inputs, labels, mask = O.prepare_batch(batch, 64, False)
    g32, l32, _ = O.compute_grads(params, cfg, inputs, labels, mask, torch.float32)
    g16, l16, _ = O.compute_grads(params, cfg, inputs, labels, mask, torch.bfloat16)
    assert abs(float(l32) - float(l16)) / float(l32) < 2e-3
    for k in g32:
        rel = float((g16[k].double() - g32[k].double()).norm() / g32[k].double().norm())
        assert rel < 5e-2, (k, rel)


def test_pmean_and_linearity_of_the_batch_mean():
    """compute_grads on a batch == pmean of compute_grads on its equal shards (what lax.pmean relies on)."""
    cfg = O.ModelConfig(vocab_size=101, seq_len=32, dim=64, expand=8 / 3, n_layers=1, n_heads=1)
    params = O.init_params(cfg, seed=4)
    batch = O.synthetic_batch(4, 32, 101, seed=11)
    inputs, labels, _ = O.prepare_batch(batch, 32, False)
    g_all, l_all, a_all = O.compute_grads(params, cfg, inputs, labels)
    parts = [O.compute_grads(params, cfg, inputs[i:i + 2], labels[i:i + 2]) for i in (0, 2)]
    g_m, l_m, a_m = O.pmean_grads(parts)
    assert abs(float(l_all) - float(l_m)) < 1e-6
    for k in g_all:
        assert torch.allclose(g_all[k], g_m[k], atol=1e-7, rtol=1e-4), k


def test_clip_and_adamw_closed_forms():
    g = {"a": torch.tensor([3.0, 4.0]), "b": torch.tensor([12.0])}  # global norm 13
    c = O.clip_grads(g, 1.0)
    s = 1.0 / (13.0 + 1e-6)
    assert torch.allclose(c["a"], g["a"] * s) and torch.allclose(c["b"], g["b"] * s)
    assert O.clip_grads(g, None) is g
    c = O.clip_grads(g, 100.0)
    assert torch.equal(c["a"], g["a"])
    p = {"w": torch.tensor([1.0])}
    gr = {"w": torch.tensor([0.5])}
    z = {"w": torch.zeros(1)}
    np_, m, v = O.adamw_step(p, gr, z, z, 1, lr=0.1, b1=0.9, b2=0.999, eps=1e-8, weight_decay=0.1)
    # first step: m_hat = g, v_hat = g^2 -> update = sign(g) + wd*p
    assert abs(float(np_["w"]) - (1.0 - 0.1 * (1.0 + 0.1))) < 1e-6


# ------------------------------------------------------------------------------------------------ golden fixtures
def _load(name):
    with open(os.path.join(GOLDEN, name + ".json")) as f:
        return json.load(f)


@pytest.mark.parametrize("name", ["small_ragged", "c1_docmask"])
def test_oracle_reproduces_golden(name):
    gold = _load(name)
    spec = gold["spec"]
    cfg = O.ModelConfig(vocab_size=spec["vocab_size"], seq_len=spec["seq_len"], dim=spec["dim"], expand=8 / 3,
                        n_layers=spec["n_layers"], n_heads=spec["n_heads"])
    params = O.init_params(cfg, seed=spec["seed"])
    assert O.count_params(params) == gold["n_params"]
    batch = O.synthetic_batch(spec["batch"], spec["seq_len"], spec["vocab_size"], seed=spec["data_seed"],
                              doc_mean=spec["doc_mean"])
    assert int(np.asarray(batch["input_ids"]).astype(np.int64).sum()) == gold["input_ids_sum"]
    assert batch.get("docs_lengths") == gold["docs_lengths"]
    use_doc = spec["doc_mean"] is not None
    inputs, labels, mask = O.prepare_batch(batch, spec["seq_len"], use_doc)
    grads, loss, acc = O.compute_grads(params, cfg, inputs, labels, mask, torch.float32)
    assert abs(float(loss) - gold["loss"]) <= 2e-6 * gold["loss"]
    assert abs(float(acc) - gold["acc"]) < 1e-7
    for k, rec in gold["grads"].items():
        g = grads[k]
        assert list(g.shape) == rec["shape"]
        assert abs(float(g.double().norm()) - rec["norm"]) <= 1e-4 * rec["norm"] + 1e-12, k
        flat = g.reshape(-1)
        for i, val in zip(rec["index"], rec["values"]):
            assert abs(float(flat[i]) - val) <= 1e-4 * abs(val) + 1e-4 * rec["norm"] / math.sqrt(flat.numel()), (k, i)
    if "ggn" in gold:  # the GGN matvec fixture along the seeded direction of tools/make_golden.py
        g = torch.Generator().manual_seed(gold["ggn"]["seed"])
        vec = {k: torch.randn(v.shape, generator=g) * (0.02 if v.dim() > 1 else 0.1) for k, v in params.items()}
        gv = O.ggn_matvec(params, cfg, vec, inputs, labels, mask, torch.float32)
        vGv = float(sum((vec[k].double() * gv[k].double()).sum() for k in gv))
        assert abs(vGv - gold["ggn"]["vGv"]) <= 1e-4 * gold["ggn"]["vGv"]
        for k, rec in gold["ggn"]["leaves"].items():
            assert abs(float(gv[k].double().norm()) - rec["norm"]) <= 1e-3 * rec["norm"] + 1e-12, k


# ------------------------------------------------------------------------------------------------ GGN matvec (N4)
def test_ce_hessian_vec_is_the_hessian_of_the_row_loss():
    """optim/ggn_utils.py:19-53: H v with H = diag(p) - p p^T equals autograd's Hessian of -log softmax(z)[y]."""
    g = torch.Generator().manual_seed(0)
    z = torch.randn(3, 7, generator=g, dtype=torch.float64)
    y = torch.tensor([1, 6, 3])
    v = torch.randn(3, 7, generator=g, dtype=torch.float64)
    hv = O.softmax_cross_entropy_hessian_vec(z, y, v)
    for r in range(3):
        H = torch.autograd.functional.hessian(lambda zz: torch.logsumexp(zz, -1) - zz[y[r]], z[r])
        assert torch.allclose(hv[r], H @ v[r], atol=1e-12)
    assert torch.allclose(hv.sum(-1), torch.zeros(3, dtype=torch.float64), atol=1e-12)  # 1^T H = 0


@pytest.mark.parametrize("use_doc", [False, True])
def test_ggn_matvec_equals_explicit_jt_h_j(use_doc):
    """make_ggn_matvec_fn_lm (optim/ggn_utils.py:69-153) on a micro model in fp64: the jvp / vjp composition equals
    J^T blockdiag(H_token) J v / (B*T) with J built explicitly by autograd; symmetric and positive semi-definite."""
    cfg = O.ModelConfig(vocab_size=11, seq_len=6, dim=8, expand=8 / 3, n_layers=1, n_heads=2)
    params = O.init_params(cfg, seed=3, dtype=torch.float64)
    params = {k: v * 10 for k, v in params.items()}  # away from the near-linear regime of the 0.02 init
    g = torch.Generator().manual_seed(4)
    ids = torch.randint(0, 11, (2, 6), generator=g)
    labels = torch.randint(0, 11, (2, 6), generator=g)
    mask = None
    if use_doc:
        mask = torch.stack([O.intra_doc_causal_mask([2, 4], 6), O.intra_doc_causal_mask([3, 1, 2], 6)])
    names = sorted(params)
    sizes = [params[k].numel() for k in names]

    def unflat(x):
        out, o = {}, 0
        for k, n in zip(names, sizes):
            out[k] = x[o:o + n].view(params[k].shape)
            o += n
        return out

    flat = torch.cat([params[k].reshape(-1) for k in names])
    J = torch.autograd.functional.jacobian(
        lambda x: O.forward(unflat(x), cfg, ids, mask, torch.float64).reshape(-1), flat)  # [B*T*V, P]
    logits = O.forward(params, cfg, ids, mask, torch.float64).reshape(12, 11)
    p = torch.softmax(logits, -1)
    H = torch.block_diag(*[torch.diag(p[r]) - torch.outer(p[r], p[r]) for r in range(12)])
    G = J.t() @ H @ J / 12.0
    v = torch.randn(flat.numel(), generator=g, dtype=torch.float64)
    gv = O.ggn_matvec(params, cfg, unflat(v), ids, labels, mask, torch.float64)
    gv_flat = torch.cat([gv[k].reshape(-1) for k in names])
    assert torch.allclose(gv_flat, G @ v, rtol=1e-8, atol=1e-10)
    assert torch.allclose(G, G.t(), atol=1e-12) and float(v @ gv_flat) >= 0
    # probe batches: the mean of the per-probe products (:137-150)
    both = O.ggn_matvec(params, cfg, unflat(v), torch.stack([ids, ids.flip(0)]), torch.stack([labels, labels]),
                        None if mask is None else torch.stack([mask, mask.flip(0)]), torch.float64)
    other = O.ggn_matvec(params, cfg, unflat(v), ids.flip(0), labels, None if mask is None else mask.flip(0),
                         torch.float64)
    for k in names:
        assert torch.allclose(both[k], 0.5 * (gv[k] + other[k]), atol=1e-12)
