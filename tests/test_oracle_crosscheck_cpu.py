"""An independent second opinion for the oracle ("Oracle cross-check"): every layer of
oracle/lm_oracle.py against an implementation its author did not write — the stock PyTorch operators
(scaled_dot_product_attention, F.rms_norm, F.cross_entropy, nn.Embedding, nn.Linear, torch.optim.AdamW,
complex multiplication for RoPE) — on random inputs in fp32 and fp64, and the whole forward / backward against an
nn.Module assembled only from those stock parts.

This cannot pin the oracle to the reference (the reference is JAX/Flax and holds no golden vectors — the oracle stays
"parity unpinned"), but it removes the single-author failure mode: a misreading of what "RMSNorm", "causal softmax
attention", "cross-entropy with integer labels", "decoupled AdamW" or "interleaved-pair RoPE" compute would have to be
shared with PyTorch's own implementations to go unnoticed.  What remains ASSUMED about flax/optax/jax semantics (and
cannot be checked without them) is listed in DESIGN.md §3.

Reference lines each check stands for: models/LM/transformer.py:41-47 (norm), :233-240 (attention), :110-134 (GLU),
:346-407 (model); models/LM/embedding.py:8-67 (RoPE); train_lm.py:181-186 (loss / accuracy); optim/factory.py:193-205
(optax.adamw)."""
import math

import pytest
import torch
import torch.nn.functional as F
from torch import nn

from oracle import lm_oracle as O


def _rel(a, b):
    a, b = a.detach(), b.detach()
    return float((a.double() - b.double()).norm() / (b.double().norm() + 1e-30))


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 2e-6), (torch.float64, 1e-13)])
@pytest.mark.parametrize("doc", [False, True])
def test_attention_against_torch_sdpa(dtype, tol, doc):
    g = torch.Generator().manual_seed(0)
    B, T, N, H = 2, 96, 3, 64
    q, k, v = (torch.randn(B, T, N, H, generator=g, dtype=dtype) for _ in range(3))
    if doc:
        batch = {"input_ids": torch.zeros(B, T + 1, dtype=torch.int32).numpy(), "docs_lengths": [[40, 30, 27], [97]]}
        mask = O.build_attn_mask(batch, T)  # bool [B, T, T], True = attend (train_lm.py:107-131)
        got = O.dot_product_attention(q, k, v, mask[:, None], is_causal=False)
        want = F.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2),
                                              attn_mask=mask[:, None])
    else:
        got = O.dot_product_attention(q, k, v, None, is_causal=True)
        want = F.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), is_causal=True)
    assert _rel(got, want.transpose(1, 2)) < tol


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-6), (torch.float64, 1e-14)])
def test_rmsnorm_against_torch_rms_norm(dtype, tol):
    g = torch.Generator().manual_seed(1)
    x = torch.randn(5, 17, 256, generator=g, dtype=dtype) * 3.0
    scale = torch.randn(256, generator=g, dtype=dtype)
    got = O.rmsnorm(x, scale, 1e-6, dtype)
    want = F.rms_norm(x, (256,), weight=scale, eps=1e-6)
    assert _rel(got, want) < tol


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-6), (torch.float64, 1e-14)])
def test_loss_and_accuracy_against_torch_cross_entropy(dtype, tol):
    g = torch.Generator().manual_seed(2)
    logits = torch.randn(3, 11, 501, generator=g, dtype=dtype) * 4.0
    labels = torch.randint(0, 501, (3, 11), generator=g)
    logits[0, 0, 7] = logits[0, 0, 400] = 99.0  # a tie: the first maximum must win (jnp.argmax)
    loss, acc = O.loss_and_acc(logits, labels.to(torch.int32))
    want = F.cross_entropy(logits.reshape(-1, 501), labels.reshape(-1), reduction="mean")
    assert abs(float(loss) - float(want)) < tol * max(1.0, abs(float(want)))
    pred = logits.argmax(-1)
    assert int(pred[0, 0]) == 7
    assert float(acc) == pytest.approx(float((pred == labels).float().mean()), abs=1e-7)


def test_rope_against_complex_multiplication():
    """apply_rotary_emb_complex_like == (x[2i] + i x[2i+1]) * exp(i t theta_i), computed here with torch's complex type."""
    g = torch.Generator().manual_seed(3)
    B, T, N, H, theta = 2, 33, 3, 64, 500000.0
    q = torch.randn(B, T, N, H, generator=g, dtype=torch.float64)
    k = torch.randn(B, T, N, H, generator=g, dtype=torch.float64)
    fc = O.precompute_freqs_cis(H, T, theta=theta)
    rq, rk = O.apply_rotary_emb_complex_like(q, k, fc)
    inv = 1.0 / (theta ** (torch.arange(0, H, 2, dtype=torch.float64) / H))
    ang = torch.outer(torch.arange(T, dtype=torch.float64), inv)
    rot = torch.polar(torch.ones_like(ang), ang)[None, :, None, :]
    for x, rx in ((q, rq), (k, rk)):
        want = torch.view_as_real(torch.view_as_complex(x.reshape(B, T, N, H // 2, 2).contiguous()) * rot).reshape(B, T, N, H)
        assert _rel(rx, want) < 3e-7  # the table itself is fp32 (embedding.py:8-26)


def test_adamw_against_torch_optim_adamw():
    """optax.adamw(lr, b1, b2, eps, weight_decay) — update = m_hat / (sqrt(v_hat) + eps) + wd * p, p -= lr * update —
    is algebraically torch.optim.AdamW (decoupled decay p *= 1 - lr wd, eps outside the bias-corrected sqrt)."""
    g = torch.Generator().manual_seed(4)
    shapes = {"a": (7, 5), "b": (13,)}
    p0 = {k: torch.randn(s, generator=g, dtype=torch.float64) for k, s in shapes.items()}
    tp = {k: nn.Parameter(v.clone()) for k, v in p0.items()}
    opt = torch.optim.AdamW(list(tp.values()), lr=3e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.1)
    params = {k: v.clone() for k, v in p0.items()}
    m = {k: torch.zeros_like(v) for k, v in p0.items()}
    v_ = {k: torch.zeros_like(v) for k, v in p0.items()}
    for step in range(1, 6):
        grads = {k: torch.randn(s, generator=g, dtype=torch.float64) for k, s in shapes.items()}
        for k in tp:
            tp[k].grad = grads[k].clone()
        opt.step()
        params, m, v_ = O.adamw_step(params, grads, m, v_, step, 3e-3, 0.9, 0.999, 1e-8, 0.1)
        for k in tp:
            assert _rel(params[k], tp[k].detach()) < 1e-12, (step, k)


def test_clip_against_torch_clip_grad_norm():
    g = torch.Generator().manual_seed(5)
    grads = {"a": torch.randn(40, 3, generator=g), "b": torch.randn(9, generator=g) * 5}
    ps = [nn.Parameter(torch.zeros_like(v)) for v in grads.values()]
    for p, v in zip(ps, grads.values()):
        p.grad = v.clone()
    torch.nn.utils.clip_grad_norm_(ps, 1.0)  # same formula: g * min(1, max_norm / (norm + 1e-6))
    got = O.clip_grads(grads, 1.0)
    for p, k in zip(ps, grads):
        assert _rel(got[k], p.grad) < 1e-6


# ------------------------------------------------------------------------------------------------ whole model
class _StockBlock(nn.Module):
    def __init__(self, d, n_heads, hid, eps, theta, seq_len):
        super().__init__()
        self.d, self.nh, self.eps = d, n_heads, eps
        self.attn_norm = nn.Parameter(torch.ones(d))
        self.mlp_norm = nn.Parameter(torch.ones(d))
        self.w_qkv = nn.Linear(d, 3 * d, bias=False)
        self.w_out = nn.Linear(d, d, bias=False)
        self.fc_gate = nn.Linear(d, hid, bias=False)
        self.fc_up = nn.Linear(d, hid, bias=False)
        self.fc2 = nn.Linear(hid, d, bias=False)
        hd = d // n_heads
        inv = 1.0 / (theta ** (torch.arange(0, hd, 2, dtype=torch.float32) / hd))
        ang = torch.outer(torch.arange(seq_len, dtype=torch.float32), inv)
        self.register_buffer("rot", torch.polar(torch.ones_like(ang), ang))

    def _rope(self, x):  # [B, T, N, H], interleaved pairs as complex numbers
        B, T, N, H = x.shape
        xc = torch.view_as_complex(x.float().reshape(B, T, N, H // 2, 2).contiguous())
        return torch.view_as_real(xc * self.rot[None, :T, None, :]).reshape(B, T, N, H).to(x.dtype)

    def forward(self, x, mask):
        B, T, d = x.shape
        h = F.rms_norm(x, (d,), weight=self.attn_norm, eps=self.eps)
        q, k, v = self.w_qkv(h).split(d, dim=-1)
        q, k, v = (t.reshape(B, T, self.nh, d // self.nh) for t in (q, k, v))
        q, k = self._rope(q), self._rope(k)
        o = F.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2),
                                           attn_mask=None if mask is None else mask[:, None], is_causal=mask is None)
        x = x + self.w_out(o.transpose(1, 2).reshape(B, T, d))
        h = F.rms_norm(x, (d,), weight=self.mlp_norm, eps=self.eps)
        return x + self.fc2(F.silu(self.fc_gate(h)) * self.fc_up(h))


class _StockLM(nn.Module):
    """The models/LM Transformer (glu, untied) written with stock torch.nn parts only."""

    def __init__(self, cfg: O.ModelConfig, params):
        super().__init__()
        self.embed = nn.Embedding(cfg.vocab_size, cfg.dim)
        self.blocks = nn.ModuleList(_StockBlock(cfg.dim, cfg.n_heads, cfg.hidden_dim, cfg.rmsnorm_eps, cfg.rope_theta,
                                                cfg.seq_len) for _ in range(cfg.n_layers))
        self.out_norm = nn.Parameter(torch.ones(cfg.dim))
        self.lm_head = nn.Linear(cfg.dim, cfg.vocab_size, bias=False)
        self.eps = cfg.rmsnorm_eps
        self.names = {}
        with torch.no_grad():  # flax kernels are (in, out); nn.Linear.weight is (out, in)
            self._bind("embed_tokens/embedding", self.embed.weight, params, False)
            for i, b in enumerate(self.blocks):
                p = f"layers_{i}"
                self._bind(f"{p}/attn_norm/RMSNorm_0/scale", b.attn_norm, params, False)
                self._bind(f"{p}/mlp_norm/RMSNorm_0/scale", b.mlp_norm, params, False)
                self._bind(f"{p}/attn/w_qkv/kernel", b.w_qkv.weight, params, True)
                self._bind(f"{p}/attn/w_out/kernel", b.w_out.weight, params, True)
                self._bind(f"{p}/mlp/fc_gate/kernel", b.fc_gate.weight, params, True)
                self._bind(f"{p}/mlp/fc_up/kernel", b.fc_up.weight, params, True)
                self._bind(f"{p}/mlp/fc2/kernel", b.fc2.weight, params, True)
            self._bind("out_norm/RMSNorm_0/scale", self.out_norm, params, False)
            self._bind("lm_head/kernel", self.lm_head.weight, params, True)

    def _bind(self, name, param, params, transposed):
        param.copy_(params[name].t() if transposed else params[name])
        self.names[name] = (param, transposed)

    def forward(self, ids, mask=None):
        x = self.embed(ids.long())
        for b in self.blocks:
            x = b(x, mask)
        return self.lm_head(F.rms_norm(x, (x.shape[-1],), weight=self.out_norm, eps=self.eps))

    def flax_grads(self):
        return {k: (p.grad.t() if tr else p.grad) for k, (p, tr) in self.names.items()}


@pytest.mark.parametrize("use_doc", [False, True])
def test_whole_model_forward_and_gradients_against_stock_torch_modules(use_doc):
    cfg = O.ModelConfig(vocab_size=211, seq_len=48, dim=128, expand=8 / 3, n_layers=2, n_heads=2)
    params = O.init_params(cfg, seed=6)
    g = torch.Generator().manual_seed(7)
    for k in params:  # non-trivial norm scales
        if k.endswith("scale"):
            params[k] = 1.0 + 0.2 * torch.randn(params[k].shape, generator=g)
    batch = O.synthetic_batch(3, 48, 211, seed=8, doc_mean=15 if use_doc else None)
    inputs, labels, mask = O.prepare_batch(batch, 48, use_doc)
    stock = _StockLM(cfg, params)
    logits = stock(inputs, mask)
    ref = O.forward(params, cfg, inputs, mask, torch.float32)
    assert _rel(ref, logits) < 5e-6
    loss = F.cross_entropy(logits.reshape(-1, 211), labels.long().reshape(-1))
    loss.backward()
    g_o, loss_o, _ = O.compute_grads(params, cfg, inputs, labels, mask, torch.float32)
    assert abs(float(loss_o) - float(loss)) < 1e-5 * float(loss)
    g_s = stock.flax_grads()
    assert set(g_s) == set(g_o)
    for k in g_o:
        assert _rel(g_o[k], g_s[k]) < 5e-5, k
