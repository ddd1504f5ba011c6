"""oracle/lm_oracle.py — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

CPU restatement (plain PyTorch, no custom kernels) of the training-step hot path of GeorgTirp/plainCV's
`models/LM` causal Transformer as driven by `train_lm.py`.  Only `tests/`, `__graft_entry__.smoke()` and
`bench.py`'s `cpu_baseline` / `--impl reference` legs may import this module; `plaincv_b200/` never does.

PARITY UNPINNED: the reference is JAX/Flax (train_lm.py:2,54-60) and jax/flax/optax are not installable in this
environment, the reference ships no tests or golden vectors (its only test.py prints a Dense output shape), and the
arithmetic of nn.Dense / nn.RMSNorm / nn.Embed / jax.nn.dot_product_attention / optax.softmax_cross_entropy lives
in third-party packages (flax 0.12.1, jax 0.9.0, optax 0.2.6 per requirements.txt) that are not vendored.  This
file restates their published semantics and follows the reference's own call sites line by line:

  models/LM/transformer.py   ModelConfig :13-26, Norm/RMSNorm :32-47, MLP :70-97, GLU :100-135,
                             MLPReluSquared :138-165, Attention :171-254, Block :260-335, Transformer :341-407
  models/LM/embedding.py     precompute_freqs_cis :8-26, apply_rotary_emb_complex_like :28-67
  models/LM/constructor.py   construct_model (transformer branch) :81-104
  train_lm.py                _trim_last_token :97-104, _build_attn_mask :107-131, _prepare_batch :134-148,
                             _clip_grads :173-178, _loss_and_acc :181-186, compute_grads :256-271
  data/datasets/data_prep_utils.py   intra_doc_causal_mask :14-43
  optim/factory.py           adamw branch :193-205 (optax.adamw, no mask -> decay on every leaf)
  optim/ggn_utils.py         softmax_cross_entropy_hessian_vec :19-53, make_ggn_matvec_fn_lm :69-153

Three numeric modes (`compute_dtype`):
  torch.float64  — for finite-difference gradient checks
  torch.float32  — "ground truth" the CUDA path is held to (BASELINE config C1 is exactly this)
  torch.bfloat16 — rounds where the Flax model rounds (after every Dense, after the norm, after RoPE, P before PV,
                   logits before the loss) with fp32 accumulation inside each contraction, to bound the expected
                   bf16-kernel-vs-fp32-oracle gap.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from fractions import Fraction
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

Params = Dict[str, torch.Tensor]


@dataclass
class ModelConfig:
    """models/LM/transformer.py:13-26 (dtype fields are passed per call here)."""

    vocab_size: int
    seq_len: int
    dim: int
    expand: float
    n_layers: int
    n_heads: int
    mlp: str = "glu"
    rmsnorm_eps: float = 1e-6
    tie_embeddings: bool = False
    rope_theta: float = 500000.0

    @property
    def hidden_dim(self) -> int:
        return int(self.expand * self.dim)  # transformer.py:296

    @property
    def head_dim(self) -> int:
        return self.dim // self.n_heads


def model_config_from_cfg(cfg) -> ModelConfig:
    """models/LM/constructor.py:81-97 — cfg is any mapping/object with the YAML keys."""
    get = (lambda k, d=None: cfg.get(k, d)) if isinstance(cfg, dict) else (lambda k, d=None: getattr(cfg, k, d))
    if get("model") != "transformer":
        raise NotImplementedError(f"Not implemented model: {get('model')}.")
    return ModelConfig(
        vocab_size=int(get("vocab_size")),
        seq_len=int(get("seq_len")),
        dim=int(get("d_model")),
        expand=float(Fraction(str(get("expand")))),
        n_layers=int(get("n_layers")),
        n_heads=int(get("n_heads")),
        mlp=get("mlp_class"),
        rmsnorm_eps=1e-6,
        tie_embeddings=bool(get("tie_embeddings", False)),
        rope_theta=float(get("rope_theta", 500000.0)),
    )


# ------------------------------------------------------------------------------------------------ parameters
def param_shapes(cfg: ModelConfig) -> Dict[str, Tuple[int, ...]]:
    """The Flax parameter tree ('/'-joined paths) with its shapes; kernels are (in, out)."""
    d, hid = cfg.dim, cfg.hidden_dim
    shapes: Dict[str, Tuple[int, ...]] = {"embed_tokens/embedding": (cfg.vocab_size, d)}
    for i in range(cfg.n_layers):
        p = f"layers_{i}"
        shapes[f"{p}/attn_norm/RMSNorm_0/scale"] = (d,)
        shapes[f"{p}/attn/w_qkv/kernel"] = (d, 3 * d)
        shapes[f"{p}/attn/w_out/kernel"] = (d, d)
        shapes[f"{p}/mlp_norm/RMSNorm_0/scale"] = (d,)
        if cfg.mlp == "glu":
            shapes[f"{p}/mlp/fc_gate/kernel"] = (d, hid)
            shapes[f"{p}/mlp/fc_up/kernel"] = (d, hid)
        elif cfg.mlp in ("mlp", "mlp_relu_sq"):
            shapes[f"{p}/mlp/fc1/kernel"] = (d, hid)
        else:
            raise ValueError(f"Unknown mlp type: {cfg.mlp}")
        shapes[f"{p}/mlp/fc2/kernel"] = (hid, d)
    shapes["out_norm/RMSNorm_0/scale"] = (d,)
    if not cfg.tie_embeddings:
        shapes["lm_head/kernel"] = (d, cfg.vocab_size)
    return shapes


def init_params(cfg: ModelConfig, seed: int = 0, dtype: torch.dtype = torch.float32) -> Params:
    """N(0, 0.02) everywhere, residual-branch outputs (w_out, fc2) N(0, 0.02/sqrt(2L)), norm scales = 1
    (transformer.py:185-187, :297-298, :359; flax RMSNorm scale_init = ones).  Seeded torch generator — the
    reference's jax PRNG stream cannot be reproduced here and does not need to be."""
    if cfg.dim % cfg.n_heads != 0:
        raise ValueError("dim must be divisible by n_heads")
    g = torch.Generator().manual_seed(seed)
    out: Params = {}
    resid_std = 0.02 / math.sqrt(2 * cfg.n_layers)
    for name, shape in param_shapes(cfg).items():
        if name.endswith("scale"):
            out[name] = torch.ones(shape, dtype=dtype)
        else:
            std = resid_std if ("w_out" in name or "fc2" in name) else 0.02
            out[name] = (torch.randn(shape, generator=g, dtype=torch.float32) * std).to(dtype)
    return out


def count_params(params: Params) -> int:
    return sum(int(p.numel()) for p in params.values())


# ------------------------------------------------------------------------------------------------ RoPE
def precompute_freqs_cis(dim: int, end: int, theta: float = 10000.0, condense_ratio: int = 1) -> torch.Tensor:
    """models/LM/embedding.py:8-26 -> (1, end, 1, dim/2, 2) fp32 with [...,0]=cos, [...,1]=sin."""
    if dim % 2 != 0:
        raise ValueError("dim must be even (pairs of real/imag).")
    inv_freqs = 1.0 / (theta ** (torch.arange(0, dim, 2, dtype=torch.float32) / dim))
    t = torch.arange(end, dtype=torch.float32) / float(condense_ratio)
    freqs = torch.outer(t, inv_freqs).to(torch.float32)
    cos = torch.cos(freqs)[None, :, None, :]
    sin = torch.sin(freqs)[None, :, None, :]
    return torch.stack([cos, sin], dim=-1)


def apply_rotary_emb_complex_like(q: torch.Tensor, k: torch.Tensor, freqs_cis: torch.Tensor):
    """models/LM/embedding.py:28-67: interleaved pairs (x[2i], x[2i+1]) rotated in fp32 (fp64 kept in fp64),
    result cast back to the input dtype."""
    qk = torch.cat([q, k], dim=2)
    b, t, two_h, d = qk.shape
    if d % 2 != 0:
        raise ValueError("head_dim must be even for complex-like RoPE.")
    work = torch.float64 if qk.dtype == torch.float64 else torch.float32
    qk_r2 = qk.reshape(b, t, two_h, d // 2, 2).to(work)
    cos = freqs_cis[..., 0].to(work)
    sin = freqs_cis[..., 1].to(work)
    a = qk_r2[..., 0]
    b_im = qk_r2[..., 1]
    rot_a = a * cos - b_im * sin
    rot_b = b_im * cos + a * sin
    rotated = torch.stack([rot_a, rot_b], dim=-1).reshape(b, t, two_h, d).to(q.dtype)
    h = q.shape[2]
    return rotated[:, :, :h, :], rotated[:, :, h:, :]


# ------------------------------------------------------------------------------------------------ layers
def _acc_dtype(compute_dtype: torch.dtype) -> torch.dtype:
    return torch.float64 if compute_dtype == torch.float64 else torch.float32


def dense(x: torch.Tensor, kernel: torch.Tensor, compute_dtype: torch.dtype) -> torch.Tensor:
    """flax.linen.Dense(use_bias=False, dtype=compute_dtype): input and (in,out) kernel are cast to the compute
    dtype, contracted with fp32 accumulation, output in the compute dtype."""
    acc = _acc_dtype(compute_dtype)
    xk = x.to(compute_dtype).to(acc)
    wk = kernel.to(compute_dtype).to(acc)
    return (xk @ wk).to(compute_dtype)


def rmsnorm(x: torch.Tensor, scale: torch.Tensor, eps: float, compute_dtype: torch.dtype) -> torch.Tensor:
    """flax.linen.RMSNorm(epsilon, dtype=compute_dtype): mean(x^2) over the last axis in fp32, y = x*rsqrt(ms+eps)*scale
    in fp32, cast to the compute dtype (transformer.py:41-47)."""
    acc = _acc_dtype(compute_dtype)
    xf = x.to(acc)
    ms = (xf * xf).mean(dim=-1, keepdim=True)
    mul = torch.rsqrt(ms + eps) * scale.to(acc)
    return (xf * mul).to(compute_dtype)


def silu(x: torch.Tensor) -> torch.Tensor:
    return x * torch.sigmoid(x)  # jax.nn.silu, evaluated in the compute dtype


def dot_product_attention(q, k, v, mask: Optional[torch.Tensor], is_causal: bool) -> torch.Tensor:
    """jax.nn.dot_product_attention, XLA path (transformer.py:233-240): q,k,v (B,T,N,H); logits BNTS in fp32 scaled by
    1/sqrt(H); masked positions replaced by -0.7*finfo.max before an fp32 softmax; probabilities cast to the value
    dtype; output (B,T,N,H) in the value dtype.  mask: bool broadcastable to (B,N,T,S), True = attend."""
    acc = torch.float64 if q.dtype == torch.float64 else torch.float32
    B, T, N, H = q.shape
    S = k.shape[1]
    logits = torch.einsum("btnh,bsnh->bnts", q.to(acc), k.to(acc))
    logits = logits * (1.0 / math.sqrt(H))
    neg = -0.7 * torch.finfo(acc).max
    if mask is not None:
        logits = torch.where(mask, logits, torch.full((), neg, dtype=acc))
    if is_causal:
        causal = torch.tril(torch.ones(T, S, dtype=torch.bool))
        logits = torch.where(causal[None, None], logits, torch.full((), neg, dtype=acc))
    probs = torch.softmax(logits, dim=-1).to(v.dtype)
    out = torch.einsum("bnts,bsnh->btnh", probs.to(acc), v.to(acc))
    return out.to(v.dtype)


def attention(params: Params, prefix: str, cfg: ModelConfig, x, freqs_cis, attn_mask, compute_dtype):
    """Attention.__call__, transformer.py:171-254."""
    d, n_heads = cfg.dim, cfg.n_heads
    head_dim = d // n_heads
    bsz, seqlen, _ = x.shape
    qkv = dense(x, params[f"{prefix}/w_qkv/kernel"], compute_dtype)
    q, k, v = torch.split(qkv, d, dim=-1)
    q = q.reshape(bsz, seqlen, n_heads, head_dim)
    k = k.reshape(bsz, seqlen, n_heads, head_dim)
    v = v.reshape(bsz, seqlen, n_heads, head_dim)
    q, k = apply_rotary_emb_complex_like(q, k, freqs_cis)
    mask = None
    is_causal = False
    if attn_mask is None:
        is_causal = True
    else:
        expanded = attn_mask[:, None, :, :]
        if expanded.dtype != torch.bool:
            raise NotImplementedError("additive masks (transformer.py:227-229) are not used by train_lm.py")
        mask = expanded
    out = dot_product_attention(q, k, v, mask, is_causal)
    out = out.reshape(bsz, seqlen, d)
    return dense(out, params[f"{prefix}/w_out/kernel"], compute_dtype)


def mlp_block(params: Params, prefix: str, cfg: ModelConfig, x, compute_dtype):
    """MLP / GLU / MLPReluSquared, transformer.py:70-165."""
    if cfg.mlp == "glu":
        gate = dense(x, params[f"{prefix}/fc_gate/kernel"], compute_dtype)
        up = dense(x, params[f"{prefix}/fc_up/kernel"], compute_dtype)
        h = silu(gate) * up
    elif cfg.mlp == "mlp":
        h = silu(dense(x, params[f"{prefix}/fc1/kernel"], compute_dtype))
    elif cfg.mlp == "mlp_relu_sq":
        h = torch.square(torch.relu(dense(x, params[f"{prefix}/fc1/kernel"], compute_dtype)))
    else:
        raise ValueError(f"Unknown mlp type: {cfg.mlp}")
    return dense(h, params[f"{prefix}/fc2/kernel"], compute_dtype)


def forward(params: Params, cfg: ModelConfig, input_ids: torch.Tensor, attn_mask: Optional[torch.Tensor] = None,
            compute_dtype: torch.dtype = torch.float32) -> torch.Tensor:
    """Transformer.__call__, transformer.py:346-407 -> logits (B, T, V) in the compute dtype."""
    if cfg.dim % cfg.n_heads != 0:
        raise ValueError("dim must be divisible by n_heads")
    head_dim = cfg.dim // cfg.n_heads
    emb = params["embed_tokens/embedding"]
    x = emb.to(compute_dtype)[input_ids.long()]  # nn.Embed: table cast to the compute dtype, then gathered
    freqs_cis = precompute_freqs_cis(head_dim, cfg.seq_len, theta=cfg.rope_theta)[:, : x.shape[1]]
    for i in range(cfg.n_layers):
        p = f"layers_{i}"
        h = rmsnorm(x, params[f"{p}/attn_norm/RMSNorm_0/scale"], cfg.rmsnorm_eps, compute_dtype)
        x = x + attention(params, f"{p}/attn", cfg, h, freqs_cis, attn_mask, compute_dtype)
        h = rmsnorm(x, params[f"{p}/mlp_norm/RMSNorm_0/scale"], cfg.rmsnorm_eps, compute_dtype)
        x = x + mlp_block(params, f"{p}/mlp", cfg, h, compute_dtype)
    x = rmsnorm(x, params["out_norm/RMSNorm_0/scale"], cfg.rmsnorm_eps, compute_dtype)
    if cfg.tie_embeddings:
        return dense(x, emb.t(), compute_dtype)  # embed.attend, transformer.py:394-396
    return dense(x, params["lm_head/kernel"], compute_dtype)


def loss_and_acc(logits: torch.Tensor, labels: torch.Tensor):
    """train_lm.py:181-186: fp32 (fp64 stays fp64) softmax cross-entropy with integer labels, mean over every token;
    accuracy = mean(first-max argmax == label)."""
    acc = torch.float64 if logits.dtype == torch.float64 else torch.float32
    lf = logits.to(acc)
    lse = torch.logsumexp(lf, dim=-1)
    picked = torch.gather(lf, -1, labels.long().unsqueeze(-1)).squeeze(-1)
    loss = (lse - picked).mean()
    accuracy = (torch.argmax(lf, dim=-1) == labels.long()).to(torch.float32).mean()
    return loss, accuracy


def compute_grads(params: Params, cfg: ModelConfig, inputs: torch.Tensor, labels: torch.Tensor,
                  attn_mask: Optional[torch.Tensor] = None, compute_dtype: torch.dtype = torch.float32):
    """Single-device compute_grads, train_lm.py:288-300 / :227-239 -> (grads, loss, acc)."""
    leaves = {k: v.detach().clone().requires_grad_(True) for k, v in params.items()}
    logits = forward(leaves, cfg, inputs, attn_mask, compute_dtype)
    loss, acc = loss_and_acc(logits, labels)
    loss.backward()
    grads = {k: v.grad.detach() for k, v in leaves.items()}
    return grads, loss.detach(), acc.detach()


def eval_step(params: Params, cfg: ModelConfig, inputs, labels, attn_mask=None, compute_dtype=torch.float32):
    """train_lm.py:302-311."""
    with torch.no_grad():
        return loss_and_acc(forward(params, cfg, inputs, attn_mask, compute_dtype), labels)


def pmean_grads(per_device: Sequence[Tuple[Params, torch.Tensor, torch.Tensor]]):
    """jax.lax.pmean over the 'data' axis (train_lm.py:268-270): leaf-wise mean of grads, loss, acc."""
    n = len(per_device)
    grads = {k: sum(g[0][k] for g in per_device) / n for k in per_device[0][0]}
    loss = sum(g[1] for g in per_device) / n
    acc = sum(g[2] for g in per_device) / n
    return grads, loss, acc


# ------------------------------------------------------------------------------------------------ batch / masks
def intra_doc_causal_mask(doc_boundaries: Sequence[int], max_seq_length: int) -> torch.Tensor:
    """data/datasets/data_prep_utils.py:14-43: block-diagonal of lower-triangular blocks, bool [T, T]."""
    if sum(doc_boundaries) != max_seq_length:
        raise ValueError("Sum of doc_boundaries does not match max_seq_length.")
    blocks = [torch.tril(torch.ones((n, n), dtype=torch.bool)) for n in doc_boundaries]
    return torch.block_diag(*blocks).to(torch.bool)


def trim_last_token(doc_boundaries: Sequence[int]) -> List[int]:
    """train_lm.py:97-104."""
    trimmed = list(doc_boundaries)
    if not trimmed:
        return trimmed
    trimmed[-1] -= 1
    if trimmed[-1] <= 0:
        trimmed.pop()
    return trimmed


def build_attn_mask(batch: dict, seq_len: int) -> torch.Tensor:
    """train_lm.py:107-131 -> bool [B, T, T]."""
    docs_lengths = batch.get("docs_lengths", None)
    if docs_lengths is None:
        raise ValueError("intra_doc_masking=True but docs_lengths not found in batch.")
    full_len = np.asarray(batch["input_ids"]).shape[1]
    if full_len != seq_len + 1:
        raise ValueError(f"Expected input_ids length {seq_len + 1} (seq_len+1) but got {full_len}.")
    masks = []
    for boundaries in docs_lengths:
        boundaries = trim_last_token(boundaries)
        if sum(boundaries) != seq_len:
            raise ValueError(f"Sum(doc_boundaries)={sum(boundaries)} != seq_len={seq_len}.")
        masks.append(intra_doc_causal_mask(boundaries, seq_len))
    return torch.stack(masks, dim=0)


def prepare_batch(batch: dict, seq_len: int, use_doc_mask: bool):
    """train_lm.py:134-148 -> (inputs, labels, attn_mask)."""
    input_ids = torch.as_tensor(np.asarray(batch["input_ids"]), dtype=torch.int32)
    if input_ids.shape[1] != seq_len + 1:
        raise ValueError(f"Expected input_ids length {seq_len + 1} (seq_len+1) but got {input_ids.shape[1]}.")
    inputs = input_ids[:, :-1]
    labels = input_ids[:, 1:]
    attn_mask = build_attn_mask(batch, seq_len) if use_doc_mask else None
    return inputs, labels, attn_mask


# ------------------------------------------------------------------------------------------------ "next" row N1
def clip_grads(grads: Params, max_norm: Optional[float]) -> Params:
    """train_lm.py:173-178 with optax.global_norm = sqrt(sum of squared leaves)."""
    if max_norm is None:
        return grads
    g_norm = torch.sqrt(sum((g.double() ** 2).sum() for g in grads.values())).float()
    scale = torch.minimum(torch.tensor(1.0), max_norm / (g_norm + 1e-6))
    return {k: g * scale for k, g in grads.items()}


def adamw_step(params: Params, grads: Params, m: Params, v: Params, step: int, lr: float, b1: float, b2: float,
               eps: float, weight_decay: float):
    """optax.adamw(lr, b1, b2, eps, weight_decay) without a mask (optim/factory.py:199-205); `step` is the 1-based
    count after this update.  Returns (new_params, new_m, new_v)."""
    new_p, new_m, new_v = {}, {}, {}
    for k, p in params.items():
        g = grads[k]
        mk = b1 * m[k] + (1 - b1) * g
        vk = b2 * v[k] + (1 - b2) * g * g
        m_hat = mk / (1 - b1 ** step)
        v_hat = vk / (1 - b2 ** step)
        upd = m_hat / (torch.sqrt(v_hat) + eps) + weight_decay * p
        new_p[k] = p - lr * upd
        new_m[k], new_v[k] = mk, vk
    return new_p, new_m, new_v


# ------------------------------------------------------------------------------------------------ synthetic data
# ------------------------------------------------------------------------------------------------ GGN matvec (N4)
def softmax_cross_entropy_hessian_vec(logits: torch.Tensor, labels: torch.Tensor, vec_logits: torch.Tensor):
    """optim/ggn_utils.py:19-53: H = diag(p) - p p^T per row, applied to vec_logits; fp32 (fp64 stays fp64)."""
    acc = torch.float64 if logits.dtype == torch.float64 else torch.float32
    probs = torch.softmax(logits.to(acc), dim=-1)
    v = vec_logits.to(acc)
    return probs * v - probs * torch.sum(probs * v, dim=-1, keepdim=True)


def ggn_matvec(params: Params, cfg: ModelConfig, vec: Params, input_ids: torch.Tensor, labels: torch.Tensor,
               attn_mask: Optional[torch.Tensor] = None, compute_dtype: torch.dtype = torch.float32) -> Params:
    """make_ggn_matvec_fn_lm, optim/ggn_utils.py:69-153: J^T H (J vec) / (B*T) with J = d logits / d params
    (jax.jvp :110-114, Hessian-vec :116-127, jax.vjp :129-131); input_ids [n_probe, B, T] averages over probes
    (:137-150)."""
    from torch.func import jvp, vjp
    if input_ids.dim() == 3:
        outs = [ggn_matvec(params, cfg, vec, input_ids[i], labels[i], None if attn_mask is None else attn_mask[i],
                           compute_dtype) for i in range(input_ids.shape[0])]
        return {k: sum(o[k] for o in outs) / float(len(outs)) for k in outs[0]}

    def logits_fn(p):
        return forward(p, cfg, input_ids, attn_mask, compute_dtype)

    vec = {k: vec[k] for k in params}  # same leaf order as the parameter tree
    logits, jvp_logits = jvp(logits_fn, (params,), (vec,))
    b, t, v = logits.shape
    hv2 = softmax_cross_entropy_hessian_vec(logits.reshape(b * t, v), labels.reshape(b * t), jvp_logits.reshape(b * t, v))
    hv_logits = (hv2.reshape(b, t, v) / float(b * t)).to(logits.dtype)
    _, vjp_fun = vjp(logits_fn, params)
    (hv_params,) = vjp_fun(hv_logits)
    return hv_params


def synthetic_batch(batch_size: int, seq_len: int, vocab_size: int, seed: int = 1234,
                    doc_mean: Optional[int] = None) -> dict:
    """The loader's batch contract (data/lm_loader.py:32-38): input_ids int [B, T+1] and, when doc_mean is given,
    docs_lengths: list[list[int]] summing to T+1 per row (shape of concat_chunck's output,
    data_prep_utils.py:117-141)."""
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(0, vocab_size, (batch_size, seq_len + 1), generator=g, dtype=torch.int64)
    batch = {"input_ids": ids.numpy().astype(np.int32)}
    if doc_mean is not None:
        rng = np.random.default_rng(seed + 4321)
        docs = []
        for _ in range(batch_size):
            lens, left = [], seq_len + 1
            while left > 0:
                n = int(min(left, max(1, round(rng.exponential(doc_mean)))))
                lens.append(n)
                left -= n
            docs.append(lens)
        batch["docs_lengths"] = docs
    return batch
