"""Thin Python wrappers over the C ABI (one function per kernel launch) and the torch.autograd.Functions built on
them.  Tensors are only used as device memory handles: every wrapper passes raw pointers + the current CUDA stream
to libplaincv_b200.so.  No wrapper has a PyTorch/CPU fallback.

Reference operations replaced (see include/plaincv_b200.h for the per-kernel citations):
  nn.Dense fwd/dgrad/wgrad, nn.Embed, nn.RMSNorm (+ the residual add of Block), silu(gate)*up,
  RoPE + jax.nn.dot_product_attention, optax softmax cross-entropy + argmax accuracy.
"""
from __future__ import annotations

import math
import os
from typing import Optional, Tuple

import torch

from . import _lib
from ._lib import PCV_EPI_BF16, PCV_EPI_F32, PCV_EPI_F32_ATOMIC, check

_NUM_SMS = 148
# LM head with the softmax partials in its epilogue + the one-pass loss kernels (default); PCV_HEAD_CE_FUSED=0 (A/B
# runs only): plain head GEMM + the stand-alone two-exponential softmax-CE kernel
_HEAD_CE_FUSED = os.environ.get("PCV_HEAD_CE_FUSED", "1") != "0"


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _p(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _req(t: torch.Tensor, dtype: torch.dtype, what: str) -> None:
    if not t.is_cuda:
        raise RuntimeError(f"{what}: plaincv_b200 kernels need CUDA tensors (there is no CPU fallback)")
    if t.dtype != dtype:
        raise TypeError(f"{what}: expected {dtype}, got {t.dtype}")


# ------------------------------------------------------------------------------------------------ raw launches
def gemm(a: torch.Tensor, b: torch.Tensor, c: torch.Tensor, M: int, N: int, K: int, lda: int, ldb: int, ldc: int,
         a_mn: bool, b_mn: bool, epilogue: int, splits: int = 1, alpha: float = 1.0) -> None:
    """C[M,N] (=|+=) alpha * sum_k A(m,k) B(n,k) on tcgen05 (pcv_gemm_bf16)."""
    lib = _lib.load()
    check(lib.pcv_gemm_bf16(a.data_ptr(), b.data_ptr(), c.data_ptr(), M, N, K, lda, ldb, ldc, int(a_mn), int(b_mn),
                            epilogue, splits, alpha, _stream()), "pcv_gemm_bf16")


def linear_fwd(x: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
    """y[M,Np] = x[M,K] @ w[K,Np]; w is the bf16 copy of a flax (in,out) kernel (row stride Np)."""
    M, K = x.shape
    Np = w.shape[1]
    y = torch.empty((M, Np), dtype=torch.bfloat16, device=x.device)
    gemm(x, w, y, M, Np, K, K, Np, Np, False, True, PCV_EPI_BF16)
    return y


def linear_fwd_rope(x: torch.Tensor, w: torch.Tensor, cos_sin: torch.Tensor, T: int, rope_cols: int) -> torch.Tensor:
    """qkv[M,3d] = rope(x @ w_qkv): the QKV projection with RoPE applied to the q | k columns in the GEMM epilogue."""
    M, K = x.shape
    Np = w.shape[1]
    y = torch.empty((M, Np), dtype=torch.bfloat16, device=x.device)
    check(_lib.load().pcv_gemm_bf16_rope(x.data_ptr(), w.data_ptr(), y.data_ptr(), M, Np, K, K, Np, Np,
                                         cos_sin.data_ptr(), T, rope_cols, _stream()), "pcv_gemm_bf16_rope")
    return y


def linear_fwd_swiglu(x: torch.Tensor, w_gu: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """(gu, h): gu[M,2*hid] = x @ [fc_gate | fc_up] and h = silu(g) * u from one GEMM launch (fused epilogue)."""
    M, K = x.shape
    two_h = w_gu.shape[1]
    hid = two_h // 2
    gu = torch.empty((M, two_h), dtype=torch.bfloat16, device=x.device)
    h = torch.empty((M, hid), dtype=torch.bfloat16, device=x.device)
    check(_lib.load().pcv_gemm_bf16_swiglu(x.data_ptr(), w_gu.data_ptr(), gu.data_ptr(), h.data_ptr(), M, hid, K, K,
                                           two_h, two_h, hid, _stream()), "pcv_gemm_bf16_swiglu")
    return gu, h


def linear_dgrad_dswiglu(dy: torch.Tensor, w_down: torch.Tensor, gu: torch.Tensor) -> torch.Tensor:
    """dgu[M,2*hid] = swiglu'(gu) applied to dh = dy[M,d] @ w_down[hid,d]^T, dh never materialised."""
    M, K = dy.shape
    hid = w_down.shape[0]
    dgu = torch.empty_like(gu)
    check(_lib.load().pcv_gemm_bf16_dswiglu(dy.data_ptr(), w_down.data_ptr(), gu.data_ptr(), dgu.data_ptr(), M, hid, K,
                                            K, w_down.shape[1], gu.shape[1], _stream()), "pcv_gemm_bf16_dswiglu")
    return dgu


def rope_qk_(qkv: torch.Tensor, cos_sin: torch.Tensor, B: int, T: int, n_heads: int) -> torch.Tensor:
    """In-place RoPE of the q | k columns of qkv (stand-alone form of the fused epilogue above)."""
    check(_lib.load().pcv_rope_qk(qkv.data_ptr(), cos_sin.data_ptr(), B, T, n_heads, _stream()), "pcv_rope_qk")
    return qkv


def linear_dgrad(dy: torch.Tensor, w: torch.Tensor, alpha: float = 1.0) -> torch.Tensor:
    """dx[M,K] = dy[M,Np] @ w[K,Np]^T."""
    M, Np = dy.shape
    K = w.shape[0]
    dx = torch.empty((M, K), dtype=torch.bfloat16, device=dy.device)
    gemm(dy, w, dx, M, K, Np, Np, Np, K, False, False, PCV_EPI_BF16, 1, alpha)
    return dx


def linear_wgrad(x: torch.Tensor, dy: torch.Tensor, dw: torch.Tensor, alpha: float = 1.0) -> None:
    """dw[K,Np] (fp32) += x[M,K]^T @ dy[M,Np]  (split-K, red.add epilogue; split factor picked by the library)."""
    M, K = x.shape
    Np = dy.shape[1]
    gemm(x, dy, dw, K, Np, M, K, Np, Np, True, True, PCV_EPI_F32_ATOMIC, 0, alpha)  # splits=0: chosen by the library


def cast_pad_bf16(src: torch.Tensor, dst: torch.Tensor, rows: int, cols: int, src_ld: int, dst_ld: int) -> None:
    lib = _lib.load()
    check(lib.pcv_cast_pad_bf16(src.data_ptr(), dst.data_ptr(), rows, cols, src_ld, dst_ld, _stream()),
          "pcv_cast_pad_bf16")


def cast_f32_bf16(src: torch.Tensor, dst: torch.Tensor) -> None:
    """dst = src converted between fp32 and bf16 (flat, same numel), direction from the dtypes."""
    to_bf16 = src.dtype == torch.float32
    _req(src, torch.float32 if to_bf16 else torch.bfloat16, "cast_f32_bf16")
    _req(dst, torch.bfloat16 if to_bf16 else torch.float32, "cast_f32_bf16")
    check(_lib.load().pcv_cast_f32_bf16(src.data_ptr(), dst.data_ptr(), src.numel(), int(to_bf16), _stream()),
          "pcv_cast_f32_bf16")


def sumsq_partials(g: torch.Tensor, partials: torch.Tensor) -> None:
    """partials[b] = this block's share of sum(g^2) (optax.global_norm inside _clip_grads, train_lm.py:173-178)."""
    _req(g, torch.float32, "sumsq_partials")
    _req(partials, torch.float32, "sumsq_partials")
    check(_lib.load().pcv_sumsq_partials(g.data_ptr(), g.numel(), partials.data_ptr(), partials.numel(), _stream()),
          "pcv_sumsq_partials")


def adamw_step(p: torch.Tensor, g: torch.Tensor, m: torch.Tensor, v: torch.Tensor, p_bf16: Optional[torch.Tensor],
               partials: Optional[torch.Tensor], norm_out: Optional[torch.Tensor], max_norm: float, lr: float, b1: float,
               b2: float, eps: float, weight_decay: float, step: int) -> None:
    """Clip + optax.adamw update of flat fp32 buffers in one pass (train_lm.py:173-178, 316-353;
    optim/factory.py:193-205); p_bf16 receives the refreshed compute copy."""
    for t in (p, g, m, v):
        _req(t, torch.float32, "adamw_step")
    check(_lib.load().pcv_adamw_step(
        p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), p_bf16.data_ptr() if p_bf16 is not None else None,
        p.numel(), partials.data_ptr() if partials is not None else None,
        partials.numel() if partials is not None else 0, norm_out.data_ptr() if norm_out is not None else None,
        float(max_norm), float(lr), float(b1), float(b2), float(eps), float(weight_decay), int(step), _stream()),
        "pcv_adamw_step")


def embed_fwd(ids: torch.Tensor, table_bf16: torch.Tensor, d: int) -> torch.Tensor:
    M = ids.numel()
    x = torch.empty((M, d), dtype=torch.bfloat16, device=ids.device)
    check(_lib.load().pcv_embed_fwd(ids.data_ptr(), table_bf16.data_ptr(), x.data_ptr(), M, d, _stream()),
          "pcv_embed_fwd")
    return x


def embed_bwd(ids: torch.Tensor, dx: torch.Tensor, dE: torch.Tensor, d: int, ldE: int) -> None:
    check(_lib.load().pcv_embed_bwd(ids.data_ptr(), dx.data_ptr(), dE.data_ptr(), ids.numel(), d, ldE, _stream()),
          "pcv_embed_bwd")


def rmsnorm_residual_fwd(res_in: torch.Tensor, delta: Optional[torch.Tensor], scale: torch.Tensor, eps: float):
    """-> (res_out, y, rstd); res_out is res_in itself when delta is None."""
    M, d = res_in.shape
    y = torch.empty_like(res_in)
    rstd = torch.empty((M,), dtype=torch.float32, device=res_in.device)
    res_out = torch.empty_like(res_in) if delta is not None else res_in
    check(_lib.load().pcv_rmsnorm_residual_fwd(res_in.data_ptr(), _p(delta), scale.data_ptr(),
                                               res_out.data_ptr() if delta is not None else None, y.data_ptr(),
                                               rstd.data_ptr(), M, d, eps, _stream()), "pcv_rmsnorm_residual_fwd")
    return res_out, y, rstd


def rmsnorm_residual_bwd(dy: torch.Tensor, dres_out: Optional[torch.Tensor], res_out: torch.Tensor,
                         scale: torch.Tensor, rstd: torch.Tensor, dscale: torch.Tensor) -> torch.Tensor:
    M, d = dy.shape
    dres_in = torch.empty_like(dy)
    check(_lib.load().pcv_rmsnorm_residual_bwd(dy.data_ptr(), _p(dres_out), res_out.data_ptr(), scale.data_ptr(),
                                               rstd.data_ptr(), dres_in.data_ptr(), dscale.data_ptr(), M, d,
                                               _stream()), "pcv_rmsnorm_residual_bwd")
    return dres_in


def swiglu_fwd(gu: torch.Tensor) -> torch.Tensor:
    M, two_h = gu.shape
    h = torch.empty((M, two_h // 2), dtype=torch.bfloat16, device=gu.device)
    check(_lib.load().pcv_swiglu_fwd(gu.data_ptr(), h.data_ptr(), M, two_h // 2, _stream()), "pcv_swiglu_fwd")
    return h


def swiglu_bwd(gu: torch.Tensor, dh: torch.Tensor) -> torch.Tensor:
    M, two_h = gu.shape
    dgu = torch.empty_like(gu)
    check(_lib.load().pcv_swiglu_bwd(gu.data_ptr(), dh.data_ptr(), dgu.data_ptr(), M, two_h // 2, _stream()),
          "pcv_swiglu_bwd")
    return dgu


_ACT_MODE = {"mlp": 0, "mlp_relu_sq": 1}


def act_fwd(x: torch.Tensor, mlp: str) -> torch.Tensor:
    """h = silu(x) (MLP, transformer.py:70-97) or relu(x)^2 (MLPReluSquared, :138-165) on a bf16 Dense output."""
    h = torch.empty_like(x)
    check(_lib.load().pcv_act_fwd(x.data_ptr(), h.data_ptr(), x.numel(), _ACT_MODE[mlp], _stream()), "pcv_act_fwd")
    return h


def act_bwd(x: torch.Tensor, dh: torch.Tensor, mlp: str) -> torch.Tensor:
    dx = torch.empty_like(x)
    check(_lib.load().pcv_act_bwd(x.data_ptr(), dh.data_ptr(), dx.data_ptr(), x.numel(), _ACT_MODE[mlp], _stream()),
          "pcv_act_bwd")
    return dx


def head_logits(hf: torch.Tensor, head) -> torch.Tensor:
    """logits[M, V_pad] of the LM head: hf @ lm_head/kernel, or hf @ embedding^T when the embeddings are tied
    (embed.attend, transformer.py:394-396: the table [V, d] is the K-major B operand as it lies)."""
    if not head.tied:
        return linear_fwd(hf, head.w_lm)
    M, d = hf.shape
    logits = torch.empty((M, head.v_pad), dtype=torch.bfloat16, device=hf.device)
    gemm(hf, head.w_lm, logits, M, head.v_pad, d, d, d, head.v_pad, False, False, PCV_EPI_BF16)
    return logits


def head_logits_ce(hf: torch.Tensor, head) -> Tuple[torch.Tensor, torch.Tensor]:
    """(logits[M, V_pad] bf16, ce_ms): the LM head with the first half of softmax cross-entropy in its epilogue
    (pcv_gemm_bf16_ce): per (row, 128-column slot) the (max, sum exp(logit - max)) of the bf16-rounded logits over the
    vocab_size real columns, float2 [ceil(V_pad / 128)][M]."""
    M, d = hf.shape
    vp = head.v_pad
    logits = torch.empty((M, vp), dtype=torch.bfloat16, device=hf.device)
    ce_ms = torch.empty(((vp + 127) // 128, M, 2), dtype=torch.float32, device=hf.device)
    ldw = d if head.tied else vp
    check(_lib.load().pcv_gemm_bf16_ce(hf.data_ptr(), head.w_lm.data_ptr(), logits.data_ptr(), M, vp, d, d, ldw, vp,
                                       0 if head.tied else 1, head.vocab, ce_ms.data_ptr(), _stream()),
          "pcv_gemm_bf16_ce")
    return logits, ce_ms


def _ce_means(lib, row_loss: torch.Tensor, row_correct: torch.Tensor, dlogits: Optional[torch.Tensor], pad) -> torch.Tensor:
    """[mean loss, mean acc] over the rows that count.  pad = None: all M rows (pcv_mean2).  pad = (B, T, T_run): the
    sequences were padded from T to T_run positions (T not a multiple of the attention tile, model.Transformer
    ._prep_padded); the pad rows are left out of both means and their d(logits) rows are zeroed, so that nothing flows
    back from them (a pad query does see real keys)."""
    M = row_loss.shape[0]
    if pad is None:
        out2 = torch.empty((2,), dtype=torch.float32, device=row_loss.device)
        check(lib.pcv_mean2(row_loss.data_ptr(), row_correct.data_ptr(), out2.data_ptr(), M, _stream()), "pcv_mean2")
        return out2
    B, T, T_run = pad
    inv = 1.0 / float(B * T)
    out2 = torch.stack([row_loss.view(B, T_run)[:, :T].sum(), row_correct.view(B, T_run)[:, :T].sum()]) * inv
    if dlogits is not None:
        dlogits.view(B, T_run, dlogits.shape[1])[:, T:, :].zero_()
    return out2


def ce_from_stats(logits: torch.Tensor, labels: torch.Tensor, ce_ms: torch.Tensor, V: int, grad_scale: float, pad=None):
    """-> (out2 = [mean loss, mean acc], dlogits): second half of the fused head + cross-entropy (pcv_ce_from_stats);
    dlogits overwrites `logits`."""
    M, ldl = logits.shape
    row_loss = torch.empty((M,), dtype=torch.float32, device=logits.device)
    row_correct = torch.empty((M,), dtype=torch.float32, device=logits.device)
    lib = _lib.load()
    check(lib.pcv_ce_from_stats(logits.data_ptr(), labels.data_ptr(), ce_ms.data_ptr(), ce_ms.shape[0], logits.data_ptr(),
                                row_loss.data_ptr(), row_correct.data_ptr(), M, V, ldl, grad_scale, _stream()),
          "pcv_ce_from_stats")
    return _ce_means(lib, row_loss, row_correct, logits, pad), logits


def head_dgrad(dlogits: torch.Tensor, head, alpha: float) -> torch.Tensor:
    if not head.tied:
        return linear_dgrad(dlogits, head.w_lm, alpha)
    M = dlogits.shape[0]
    d = head.w_lm.shape[1]
    dx = torch.empty((M, d), dtype=torch.bfloat16, device=dlogits.device)
    # dx[M, d] = dlogits[M, V] @ E[V, d]: B(n = d, k = V) is the table read MN-major
    gemm(dlogits, head.w_lm, dx, M, d, head.v_pad, head.v_pad, d, d, False, True, PCV_EPI_BF16, 1, alpha)
    return dx


def head_wgrad(hf: torch.Tensor, dlogits: torch.Tensor, head, alpha: float) -> None:
    if not head.tied:
        linear_wgrad(hf, dlogits, head.g_lm, alpha)
        return
    # dE[V, d] += dlogits[M, V]^T @ hf[M, d]  (added to the gather's scatter-add in the same fp32 rows)
    M, d = hf.shape
    gemm(dlogits, hf, head.g_lm, head.v_pad, d, M, head.v_pad, d, d, True, True, PCV_EPI_F32_ATOMIC, 0, alpha)


def attn_fwd(qkv: torch.Tensor, doc_start: Optional[torch.Tensor], B: int, T: int,
             n_heads: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """qkv: q and k columns already rotated (linear_fwd_rope / rope_qk_)."""
    d = n_heads * 64
    out = torch.empty((B * T, d), dtype=torch.bfloat16, device=qkv.device)
    lse = torch.empty((B, n_heads, T), dtype=torch.float32, device=qkv.device)
    check(_lib.load().pcv_attn_fwd(qkv.data_ptr(), _p(doc_start), out.data_ptr(), lse.data_ptr(),
                                   B, T, n_heads, _stream()), "pcv_attn_fwd")
    return out, lse


def attn_bwd(qkv: torch.Tensor, cos_sin: torch.Tensor, doc_start: Optional[torch.Tensor], out: torch.Tensor,
             dout: torch.Tensor, lse: torch.Tensor, B: int, T: int, n_heads: int) -> torch.Tensor:
    """qkv: rotated (as given to attn_fwd); returns the gradient w.r.t. the un-rotated qkv."""
    d = n_heads * 64
    dqkv = torch.empty_like(qkv)
    delta = torch.empty((B, n_heads, T), dtype=torch.float32, device=qkv.device)
    dq_acc = torch.empty((B * T, d), dtype=torch.float32, device=qkv.device)
    check(_lib.load().pcv_attn_bwd(qkv.data_ptr(), cos_sin.data_ptr(), _p(doc_start), out.data_ptr(), dout.data_ptr(),
                                   lse.data_ptr(), delta.data_ptr(), dq_acc.data_ptr(), dqkv.data_ptr(), B, T, n_heads,
                                   _stream()), "pcv_attn_bwd")
    return dqkv


def softmax_ce(logits: torch.Tensor, labels: torch.Tensor, V: int, grad_scale: float, write_grad: bool, pad=None):
    """-> (out2 = [mean loss, mean acc] fp32 device tensor, dlogits or None).  dlogits overwrites `logits`."""
    M, ldl = logits.shape
    row_loss = torch.empty((M,), dtype=torch.float32, device=logits.device)
    row_correct = torch.empty((M,), dtype=torch.float32, device=logits.device)
    lib = _lib.load()
    check(lib.pcv_softmax_ce(logits.data_ptr(), labels.data_ptr(), logits.data_ptr() if write_grad else None,
                             row_loss.data_ptr(), row_correct.data_ptr(), M, V, ldl, grad_scale, _stream()),
          "pcv_softmax_ce")
    return _ce_means(lib, row_loss, row_correct, logits if write_grad else None, pad), (logits if write_grad else None)


# ------------------------------------------------------------------------------------------------ tangents (N4)
def add_bf16(a: torch.Tensor, b: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    out = torch.empty_like(a) if out is None else out
    check(_lib.load().pcv_add_bf16(a.data_ptr(), b.data_ptr(), out.data_ptr(), a.numel(), _stream()), "pcv_add_bf16")
    return out


def linear_jvp(tx: torch.Tensor, w: torch.Tensor, x: torch.Tensor, tw: torch.Tensor) -> torch.Tensor:
    """Tangent of y = x @ w: ty = tx @ w + x @ tw, each product rounded to bf16 like the primal Dense output
    (jax.jvp of nn.Dense in the compute dtype), summed by pcv_add_bf16."""
    a = linear_fwd(tx, w)
    return add_bf16(a, linear_fwd(x, tw), a)


def rmsnorm_residual_jvp(x: torch.Tensor, rstd: torch.Tensor, scale: torch.Tensor, tscale: torch.Tensor,
                         tres_in: torch.Tensor, tdelta: Optional[torch.Tensor]):
    """-> (tres_out, ty) for the primal (res_out = x, rstd) of rmsnorm_residual_fwd; tres_out is tres_in when
    tdelta is None."""
    M, d = x.shape
    ty = torch.empty_like(x)
    tres_out = torch.empty_like(x) if tdelta is not None else tres_in
    check(_lib.load().pcv_rmsnorm_residual_jvp(x.data_ptr(), rstd.data_ptr(), scale.data_ptr(), tscale.data_ptr(),
                                               tres_in.data_ptr(), _p(tdelta),
                                               tres_out.data_ptr() if tdelta is not None else None, ty.data_ptr(),
                                               M, d, _stream()), "pcv_rmsnorm_residual_jvp")
    return tres_out, ty


def swiglu_jvp(gu: torch.Tensor, tgu: torch.Tensor) -> torch.Tensor:
    M, two_h = gu.shape
    th = torch.empty((M, two_h // 2), dtype=torch.bfloat16, device=gu.device)
    check(_lib.load().pcv_swiglu_jvp(gu.data_ptr(), tgu.data_ptr(), th.data_ptr(), M, two_h // 2, _stream()),
          "pcv_swiglu_jvp")
    return th


def act_jvp(x: torch.Tensor, tx: torch.Tensor, mlp: str) -> torch.Tensor:
    th = torch.empty_like(x)
    check(_lib.load().pcv_act_jvp(x.data_ptr(), tx.data_ptr(), th.data_ptr(), x.numel(), _ACT_MODE[mlp], _stream()),
          "pcv_act_jvp")
    return th


def attn_jvp(qkv: torch.Tensor, tqkv: torch.Tensor, doc_start: Optional[torch.Tensor], out: torch.Tensor,
             lse: torch.Tensor, B: int, T: int, n_heads: int) -> torch.Tensor:
    """Tangent of attn_fwd: qkv / tqkv rotated (q, k columns), out / lse from attn_fwd."""
    tout = torch.empty_like(out)
    check(_lib.load().pcv_attn_jvp(qkv.data_ptr(), tqkv.data_ptr(), _p(doc_start), out.data_ptr(), lse.data_ptr(),
                                   tout.data_ptr(), B, T, n_heads, _stream()), "pcv_attn_jvp")
    return tout


def ce_hessian_vec(logits: torch.Tensor, vec: torch.Tensor, V: int, out_scale: float,
                   out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """(diag(p) - p p^T) vec * out_scale per row of bf16 logits[M, ldl] (optim/ggn_utils.py:19-53); `out` may be vec."""
    _req(logits, torch.bfloat16, "ce_hessian_vec")
    _req(vec, torch.bfloat16, "ce_hessian_vec")
    M, ldl = logits.shape
    out = torch.empty_like(vec) if out is None else out
    check(_lib.load().pcv_ce_hessian_vec(logits.data_ptr(), vec.data_ptr(), out.data_ptr(), M, V, ldl, float(out_scale),
                                         _stream()), "pcv_ce_hessian_vec")
    return out


def rope_table(head_dim: int, seq_len: int, theta: float, device) -> torch.Tensor:
    """cos/sin table [T][hd/2][2] fp32, computed exactly as precompute_freqs_cis (models/LM/embedding.py:8-26) does:
    a one-off host-side constant, uploaded once."""
    if head_dim % 2 != 0:
        raise ValueError("dim must be even (pairs of real/imag).")
    inv_freqs = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.float32) / head_dim))
    t = torch.arange(seq_len, dtype=torch.float32)
    freqs = torch.outer(t, inv_freqs).to(torch.float32)
    return torch.stack([torch.cos(freqs), torch.sin(freqs)], dim=-1).contiguous().to(device)


# ------------------------------------------------------------------------------------------------ side stream
class _WgradStream:
    """Weight-gradient GEMMs feed nothing downstream in a block's backward, so they are launched on a second stream:
    when the dgrad GEMM on the main stream runs its last, partial wave of tiles, the idle SMs pick up the wgrad
    kernel's CTAs instead of waiting (both kernels are persistent, one CTA per SM).  `launch` orders the wgrad after
    everything already queued on the main stream; `join` makes the main stream wait for all of them (called before
    the gradients are handed to the reducer / the caller).  Tensors passed to `launch` are kept alive until `join`."""

    def __init__(self):
        self.streams = {}
        self.keep = []
        self.enabled = os.environ.get("PCV_WGRAD_STREAM", "1") != "0"  # 0: same-stream launches (A/B measurements)

    def _side(self, dev) -> "torch.cuda.Stream":
        st = self.streams.get(dev)
        if st is None:
            st = self.streams[dev] = torch.cuda.Stream(device=dev)
        return st

    def launch(self, x: torch.Tensor, dy: torch.Tensor, dw: torch.Tensor, alpha: float = 1.0) -> None:
        if not self.enabled:
            linear_wgrad(x, dy, dw, alpha)
            return
        side = self._side(x.device)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            linear_wgrad(x, dy, dw, alpha)
        self.keep.append((x, dy))

    def join(self, dev) -> None:
        if self.keep:
            torch.cuda.current_stream().wait_stream(self._side(dev))
            self.keep.clear()


wgrad_stream = _WgradStream()


# ------------------------------------------------------------------------------------------------ autograd: per op
class LinearFn(torch.autograd.Function):
    """y = x @ W for a flax (in,out) kernel.  `w_bf16` is the padded bf16 copy [K, Np]; the weight gradient is
    accumulated straight into `w_grad` (fp32 [K, Np], the parameter's slot of the flat gradient buffer)."""

    @staticmethod
    def forward(ctx, x, w_bf16, w_grad):
        ctx.save_for_backward(x, w_bf16)
        ctx.w_grad = w_grad
        return linear_fwd(x, w_bf16)

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        dy = dy.contiguous()
        dx = linear_dgrad(dy, w)
        if ctx.w_grad is not None:
            linear_wgrad(x, dy, ctx.w_grad)
        return dx, None, None


class RMSNormFn(torch.autograd.Function):
    """y = rmsnorm(x) * scale (fp32 statistics, bf16 out); dscale accumulates into `scale_grad`."""

    @staticmethod
    def forward(ctx, x, scale, scale_grad, eps):
        _, y, rstd = rmsnorm_residual_fwd(x, None, scale, eps)
        ctx.save_for_backward(x, scale, rstd)
        ctx.scale_grad = scale_grad
        return y

    @staticmethod
    def backward(ctx, dy):
        x, scale, rstd = ctx.saved_tensors
        return rmsnorm_residual_bwd(dy.contiguous(), None, x, scale, rstd, ctx.scale_grad), None, None, None


class RMSNormResidualFn(torch.autograd.Function):
    """(res_out, y) = (res + delta, rmsnorm(res + delta) * scale); dscale accumulates into `scale_grad`."""

    @staticmethod
    def forward(ctx, res, delta, scale, scale_grad, eps):
        res_out, y, rstd = rmsnorm_residual_fwd(res, delta, scale, eps)
        ctx.save_for_backward(res_out, scale, rstd)
        ctx.scale_grad = scale_grad
        return res_out, y

    @staticmethod
    def backward(ctx, dres_out, dy):
        res_out, scale, rstd = ctx.saved_tensors
        dres_in = rmsnorm_residual_bwd(dy.contiguous(), None if dres_out is None else dres_out.contiguous(), res_out,
                                       scale, rstd, ctx.scale_grad)
        return dres_in, dres_in, None, None, None


class SwiGLUFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, gu):
        ctx.save_for_backward(gu)
        return swiglu_fwd(gu)

    @staticmethod
    def backward(ctx, dh):
        (gu,) = ctx.saved_tensors
        return swiglu_bwd(gu, dh.contiguous())


class AttentionFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, cos_sin, doc_start, B, T, n_heads):
        qkv = rope_qk_(qkv.clone(), cos_sin, B, T, n_heads)  # un-fused form; BlockFn fuses RoPE into the QKV GEMM
        out, lse = attn_fwd(qkv, doc_start, B, T, n_heads)
        ctx.save_for_backward(qkv, out, lse, cos_sin, doc_start if doc_start is not None else torch.empty(0))
        ctx.has_doc = doc_start is not None
        ctx.dims = (B, T, n_heads)
        return out

    @staticmethod
    def backward(ctx, dout):
        qkv, out, lse, cos_sin, doc_start = ctx.saved_tensors
        B, T, n_heads = ctx.dims
        dqkv = attn_bwd(qkv, cos_sin, doc_start if ctx.has_doc else None, out, dout.contiguous(), lse, B, T, n_heads)
        return dqkv, None, None, None, None, None


class EmbedFn(torch.autograd.Function):
    """x = bf16(E)[ids].  `table` (the fp32 master parameter) only anchors the autograd graph; the gradient is
    scatter-added into `table_grad` directly."""

    @staticmethod
    def forward(ctx, table, ids, table_bf16, table_grad, d, ldE):
        ctx.save_for_backward(ids)
        ctx.table_grad, ctx.d, ctx.ldE = table_grad, d, ldE
        return embed_fwd(ids, table_bf16, d)

    @staticmethod
    def backward(ctx, dx):
        (ids,) = ctx.saved_tensors
        embed_bwd(ids, dx.contiguous(), ctx.table_grad, ctx.d, ctx.ldE)
        return None, None, None, None, None, None


# ------------------------------------------------------------------------------------------------ autograd: fused
class BlockFn(torch.autograd.Function):
    """One pre-norm decoder block (models/LM/transformer.py:260-335), hand-scheduled forward and backward.

    Inputs (res, delta): the residual stream entering the block is res + delta (delta = previous block's MLP output,
    None for the first block); the add is fused into the first RMSNorm.  Outputs (res2, m): residual stream after
    the attention branch and the MLP branch output, to be added by the next block / the head.  `layer` carries the
    bf16 weight copies, the fp32 norm scales and the gradient slots (see model._LayerHandles)."""

    @staticmethod
    def forward(ctx, res, delta, layer, rt):
        eps = rt.eps
        res1, h1, rstd1 = rmsnorm_residual_fwd(res, delta, layer.attn_norm, eps)
        qkv = linear_fwd_rope(h1, layer.w_qkv, rt.cos_sin, rt.T, 2 * rt.n_heads * 64)  # q, k stored rotated
        o, lse = attn_fwd(qkv, rt.doc_start, rt.B, rt.T, rt.n_heads)
        a = linear_fwd(o, layer.w_out)
        res2, h2, rstd2 = rmsnorm_residual_fwd(res1, a, layer.mlp_norm, eps)
        del a
        if layer.mlp == "glu":
            gu, hh = linear_fwd_swiglu(h2, layer.w_gu)
        else:  # MLP / MLPReluSquared: gu holds fc1's output alone
            gu = linear_fwd(h2, layer.w_gu)
            hh = act_fwd(gu, layer.mlp)
        m = linear_fwd(hh, layer.w_down)
        ctx.save_for_backward(res1, rstd1, h1, qkv, o, lse, res2, rstd2, h2, gu, hh)
        if rt.tape is not None:
            rt.tape.append((res1, rstd1, h1, qkv, o, lse, res2, rstd2, h2, gu, hh))
        ctx.layer, ctx.rt, ctx.has_delta = layer, rt, delta is not None
        return res2, m

    @staticmethod
    def backward(ctx, d_res2, d_m):
        res1, rstd1, h1, qkv, o, lse, res2, rstd2, h2, gu, hh = ctx.saved_tensors
        layer, rt = ctx.layer, ctx.rt
        d_m = d_m.contiguous()
        # MLP branch
        wg = wgrad_stream
        wg.launch(hh, d_m, layer.g_down)
        if layer.mlp == "glu":
            d_gu = linear_dgrad_dswiglu(d_m, layer.w_down, gu)
        else:
            d_gu = act_bwd(gu, linear_dgrad(d_m, layer.w_down), layer.mlp)
        wg.launch(h2, d_gu, layer.g_gu)
        d_h2 = linear_dgrad(d_gu, layer.w_gu)
        del d_gu
        d_r2 = rmsnorm_residual_bwd(d_h2, d_res2, res2, layer.mlp_norm, rstd2, layer.g_mlp_norm)
        del d_h2
        # attention branch
        wg.launch(o, d_r2, layer.g_out)
        d_o = linear_dgrad(d_r2, layer.w_out)
        d_qkv = attn_bwd(qkv, rt.cos_sin, rt.doc_start, o, d_o, lse, rt.B, rt.T, rt.n_heads)
        del d_o
        wg.launch(h1, d_qkv, layer.g_qkv)
        d_h1 = linear_dgrad(d_qkv, layer.w_qkv)
        del d_qkv
        d_r1 = rmsnorm_residual_bwd(d_h1, d_r2, res1, layer.attn_norm, rstd1, layer.g_attn_norm)
        wg.join(d_r1.device)
        if layer.on_grads_ready is not None:
            layer.on_grads_ready()
        return d_r1, (d_r1 if ctx.has_delta else None), None, None


class HeadLossFn(torch.autograd.Function):
    """out_norm -> lm_head -> softmax cross-entropy + accuracy (transformer.py:384-405, train_lm.py:181-186).
    Returns a 2-vector [loss, acc].  d(logits) is produced in the forward pass, in place of the bf16 logits, so the
    fp32 [M,V] logits of the reference are never materialised."""

    @staticmethod
    def forward(ctx, res, delta, labels, head, rt, loss_scale, unit_grad, pad=None):
        res_f, hf, rstd = rmsnorm_residual_fwd(res, delta, head.out_norm, rt.eps)
        M = hf.shape[0]
        n_tok = M if pad is None else pad[0] * pad[1]  # the mean of _loss_and_acc runs over the real tokens
        need_grad = bool(ctx.needs_input_grad[0])
        if need_grad and _HEAD_CE_FUSED:
            # the head GEMM's epilogue leaves the per-slot softmax partials, one streaming pass turns the logits into
            # d(logits) (one exponential per logit): loss kernels 0.78 -> 0.63 ms, head GEMM +0.06 ms at C3
            logits, ce_ms = head_logits_ce(hf, head)
            out2, dlogits = ce_from_stats(logits, labels, ce_ms, head.vocab, loss_scale / n_tok, pad)
        else:
            logits = head_logits(hf, head)
            out2, dlogits = softmax_ce(logits, labels, head.vocab, loss_scale / n_tok, need_grad, pad)
        if need_grad:
            ctx.save_for_backward(res_f, rstd, hf, dlogits)
        ctx.head, ctx.has_delta, ctx.unit_grad = head, delta is not None, unit_grad
        return out2

    @staticmethod
    def backward(ctx, g_out):
        res_f, rstd, hf, dlogits = ctx.saved_tensors
        head = ctx.head
        alpha = 1.0 if ctx.unit_grad else float(g_out[0])
        if head.tied:
            head_wgrad(hf, dlogits, head, alpha)
        else:
            wgrad_stream.launch(hf, dlogits, head.g_lm, alpha)
        d_hf = head_dgrad(dlogits, head, alpha)
        d_r = rmsnorm_residual_bwd(d_hf, None, res_f, head.out_norm, rstd, head.g_out_norm)
        wgrad_stream.join(d_r.device)
        if head.on_grads_ready is not None:
            head.on_grads_ready()
        return d_r, (d_r if ctx.has_delta else None), None, None, None, None, None, None
