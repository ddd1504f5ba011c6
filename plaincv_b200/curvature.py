"""Generalised Gauss-Newton matvec for the LM: the contract of optim/ggn_utils.py.

Reference (optim/ggn_utils.py):
  * softmax_cross_entropy_hessian_vec(logits, labels, vec_logits) :19-53 — (diag(p) - p p^T) v per row, fp32 softmax;
  * make_ggn_matvec_fn_lm(model_def, curvature_batch, batch_stats=None) :69-153 — returns
    ggn_matvec(params, vec_pytree, rng_key) = J^T H (J vec) / (B*T), with J from jax.jvp / jax.vjp around
    model.apply(..., attn_mask=...) and a mean over probe batches when input_ids is [n_probe, B, T].

Here J vec is a hand-scheduled tangent pass over the same sm_100a kernels as the training step (every Dense tangent is
two tcgen05 GEMMs; RMSNorm, SwiGLU and attention tangents are the kernels of csrc/curvature.cu), H is
pcv_ce_hessian_vec, and J^T is the existing backward path fed with H (J vec) as the logits cotangent.  No CPU path.

The matvec uses the module's flat gradient buffer as its accumulator (exactly like a training backward) and returns a
copy of the result: call it between optimizer steps, not between compute_grads and apply_grads.
"""
from __future__ import annotations

from typing import Dict, Optional, Sequence, Tuple, Union

import torch

from . import ops
from .model import Transformer

VecTree = Dict[str, torch.Tensor]


def softmax_cross_entropy_hessian_vec(logits: torch.Tensor, labels: Optional[torch.Tensor],
                                      vec_logits: torch.Tensor) -> torch.Tensor:
    """optim/ggn_utils.py:19-53 on bf16 CUDA logits [..., V]: H(logits) @ vec_logits per row, H = diag(p) - p p^T
    (the labels do not enter the Hessian; the argument is kept for signature parity)."""
    if logits.shape != vec_logits.shape:
        raise ValueError("logits and vec_logits must have the same shape")
    V = logits.shape[-1]
    if V % 8 != 0:
        raise ValueError("the last dimension must be a multiple of 8 (16-byte rows); pad the logits")
    l2 = logits.contiguous().view(-1, V)
    v2 = vec_logits.contiguous().view(-1, V)
    return ops.ce_hessian_vec(l2, v2, V, 1.0).view(logits.shape)


class _HeadCotangentFn(torch.autograd.Function):
    """out_norm (+ the LM head in backward) whose backward takes the logits cotangent from `box["cot"]` (bf16
    [M, V_pad]) instead of a loss: the J^T leg of the GGN product.  Forward leaves (res_f, rstd, hf) in `box["head"]`
    for the tangent pass and returns a dummy scalar to call .backward() on."""

    @staticmethod
    def forward(ctx, res, delta, head, rt, box):
        res_f, hf, rstd = ops.rmsnorm_residual_fwd(res, delta, head.out_norm, rt.eps)
        ctx.save_for_backward(res_f, rstd, hf)
        ctx.head, ctx.has_delta, ctx.box = head, delta is not None, box
        box["head"] = (res_f, rstd, hf)
        return torch.zeros((), dtype=torch.float32, device=res.device)

    @staticmethod
    def backward(ctx, g_out):
        res_f, rstd, hf = ctx.saved_tensors
        head = ctx.head
        cot = ctx.box["cot"]
        if head.tied:
            ops.head_wgrad(hf, cot, head, 1.0)
        else:
            ops.wgrad_stream.launch(hf, cot, head.g_lm, 1.0)
        d_hf = ops.head_dgrad(cot, head, 1.0)
        d_r = ops.rmsnorm_residual_bwd(d_hf, None, res_f, head.out_norm, rstd, head.g_out_norm)
        ops.wgrad_stream.join(d_r.device)
        return d_r, (d_r if ctx.has_delta else None), None, None, None


class TangentBuffers:
    """A parameter-shaped direction in the module's flat layout: fp32 (norm scales are used in fp32, as nn.RMSNorm
    does) and its bf16 cast (what promote_dtype makes of a kernel's tangent).  Pad lanes stay zero."""

    def __init__(self, model: Transformer):
        self.model = model
        n = model.layout.total
        dev = model.flat_master.device
        self.f32 = torch.zeros(n, dtype=torch.float32, device=dev)
        self.bf16 = torch.zeros(n, dtype=torch.bfloat16, device=dev)

    @torch.no_grad()
    def load(self, vec: Union[VecTree, torch.Tensor]) -> None:
        L = self.model.layout
        if isinstance(vec, torch.Tensor):
            if vec.numel() != L.total:
                raise ValueError(f"flat direction must have {L.total} elements (the module's padded layout)")
            self.f32.copy_(vec.to(device=self.f32.device, dtype=torch.float32).view(-1))
        else:
            names = {v.name for v in L.views}
            missing, extra = names - set(vec), set(vec) - names
            if missing or extra:
                raise KeyError(f"direction tree mismatch: missing={sorted(missing)[:4]} unexpected={sorted(extra)[:4]}")
            self.f32.zero_()
            for v in L.views:
                src = torch.as_tensor(vec[v.name])
                if tuple(src.shape) != tuple(v.shape):
                    raise ValueError(f"{v.name}: expected shape {tuple(v.shape)}, got {tuple(src.shape)}")
                L.param_view(self.f32, v).copy_(src.to(device=self.f32.device, dtype=torch.float32))
        ops.cast_pad_bf16(self.f32, self.bf16, 1, L.total, L.total, L.total)


def _head_matmul(hf: torch.Tensor, w: torch.Tensor, head) -> torch.Tensor:
    if not head.tied:
        return ops.linear_fwd(hf, w)
    M, d = hf.shape
    logits = torch.empty((M, head.v_pad), dtype=torch.bfloat16, device=hf.device)
    ops.gemm(hf, w, logits, M, head.v_pad, d, d, d, head.v_pad, False, False, ops.PCV_EPI_BF16)
    return logits


@torch.no_grad()
def _tangent_pass(model: Transformer, ids: torch.Tensor, rt, tape, head_state, tangent: TangentBuffers) -> torch.Tensor:
    """J vec -> bf16 [B*T, V_pad] from the primal activations of one forward pass: `tape[i]` = what block i saved
    for its backward (ops.BlockFn), `head_state` = (res_f, rstd, hf) of the final norm."""
    L, nh, B, T = model.layout, model.cfg.n_heads, rt.B, rt.T
    tf32, tb16 = tangent.f32, tangent.bf16
    tres = ops.embed_fwd(ids, L.slot_view(tb16, "embed"), L.d)
    tdelta = None
    for i, layer in enumerate(model.layers):
        p = f"layers_{i}"
        res1, rstd1, h1, qkv, o, lse, res2, rstd2, h2, gu, hh = tape[i]
        tres1, th1 = ops.rmsnorm_residual_jvp(res1, rstd1, layer.attn_norm,
                                              L.slot_view(tf32, f"{p}.attn_norm").view(-1), tres, tdelta)
        tqkv = ops.linear_jvp(th1, layer.w_qkv, h1, L.slot_view(tb16, f"{p}.w_qkv"))
        ops.rope_qk_(tqkv, rt.cos_sin, B, T, nh)  # RoPE is linear: the tangent rotates like the primal
        to = ops.attn_jvp(qkv, tqkv, rt.doc_start, o, lse, B, T, nh)
        ta = ops.linear_jvp(to, layer.w_out, o, L.slot_view(tb16, f"{p}.w_out"))
        tres2, th2 = ops.rmsnorm_residual_jvp(res2, rstd2, layer.mlp_norm,
                                              L.slot_view(tf32, f"{p}.mlp_norm").view(-1), tres1, ta)
        tgu = ops.linear_jvp(th2, layer.w_gu, h2, L.slot_view(tb16, f"{p}.w_gu"))
        thh = ops.swiglu_jvp(gu, tgu) if layer.mlp == "glu" else ops.act_jvp(gu, tgu, layer.mlp)
        tdelta = ops.linear_jvp(thh, layer.w_down, hh, L.slot_view(tb16, f"{p}.w_down"))
        tres = tres2
    head = model.head
    res_f, rstd, hf = head_state
    _, thf = ops.rmsnorm_residual_jvp(res_f, rstd, head.out_norm, L.slot_view(tf32, "out_norm").view(-1), tres, tdelta)
    t_head_w = L.slot_view(tb16, "embed" if head.tied else "w_lm")
    a = _head_matmul(thf, head.w_lm, head)
    return ops.add_bf16(a, _head_matmul(hf, t_head_w, head), a)


@torch.no_grad()
def jvp_logits(model: Transformer, input_ids: torch.Tensor, tangent: TangentBuffers, *,
               doc_start: Optional[torch.Tensor] = None, doc_lengths=None,
               attn_mask: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """jax.jvp(logits_fn, (params,), (vec,)) of optim/ggn_utils.py:110-114 -> (logits, J vec), both bf16
    [B*T, V_pad] (pad columns zero)."""
    ids, B, T, T_run, doc_start = model._prep_padded(input_ids, attn_mask, doc_lengths, doc_start)
    rt = model._runtime(B, T_run, doc_start)
    rt.tape = []
    res, delta = model._trunk(ids, rt)
    head = model.head
    res_f, hf, rstd = ops.rmsnorm_residual_fwd(res, delta, head.out_norm, rt.eps)
    logits = _head_matmul(hf, head.w_lm, head)
    tlogits = _tangent_pass(model, ids, rt, rt.tape, (res_f, rstd, hf), tangent)
    if T_run != T:  # a ragged T ran right-padded (Transformer._prep_padded): drop the pad rows
        Vp = logits.shape[1]
        logits = logits.view(B, T_run, Vp)[:, :T].reshape(B * T, Vp)
        tlogits = tlogits.view(B, T_run, Vp)[:, :T].reshape(B * T, Vp)
    return logits, tlogits


def ggn_product_accumulate(model: Transformer, input_ids: torch.Tensor, tangent: TangentBuffers, out_scale: float, *,
                           doc_start: Optional[torch.Tensor] = None, doc_lengths=None,
                           attn_mask: Optional[torch.Tensor] = None) -> None:
    """model.flat_grad += out_scale * J^T H J vec for one batch (optim/ggn_utils.py:94-131): ONE primal forward feeds
    both the tangent pass (jax.jvp, :110-114) and the transpose (jax.vjp, :129-131) — the blocks leave their saved
    activations on a tape; the Hessian-vector product (:116-127) overwrites J vec in place."""
    ids, B, T, T_run, doc_start = model._prep_padded(input_ids, attn_mask, doc_lengths, doc_start)
    rt = model._runtime(B, T_run, doc_start)
    rt.tape = []
    box: dict = {}
    with torch.enable_grad():
        res, delta = model._trunk(ids, rt)
        dummy = _HeadCotangentFn.apply(res, delta, model.head, rt, box)
    with torch.no_grad():
        head_state = box.pop("head")
        tlogits = _tangent_pass(model, ids, rt, rt.tape, head_state, tangent)
        rt.tape = None
        logits = _head_matmul(head_state[2], model.head.w_lm, model.head)
        box["cot"] = ops.ce_hessian_vec(logits, tlogits, model.cfg.vocab_size, out_scale, out=tlogits)
        if T_run != T:  # pad rows of a ragged T carry no loss: their cotangent is zero (a pad query does see real keys)
            box["cot"].view(B, T_run, -1)[:, T:, :].zero_()
        del logits, tlogits, head_state
    dummy.backward()


def vjp_logits_accumulate(model: Transformer, input_ids: torch.Tensor, cotangent: torch.Tensor, *,
                          doc_start: Optional[torch.Tensor] = None, doc_lengths=None,
                          attn_mask: Optional[torch.Tensor] = None) -> None:
    """jax.vjp(logits_fn, params)[1](cotangent) of optim/ggn_utils.py:129-130, ADDED into model.flat_grad
    (param.grad are views of it).  cotangent: bf16 [B*T, V_pad]."""
    ids, B, T, T_run, doc_start = model._prep_padded(input_ids, attn_mask, doc_lengths, doc_start)
    if tuple(cotangent.shape) != (B * T, model.layout.V_pad) or cotangent.dtype != torch.bfloat16:
        raise ValueError("cotangent must be bf16 [B*T, V_pad]")
    if T_run != T:
        cotangent = torch.nn.functional.pad(cotangent.view(B, T, -1), (0, 0, 0, T_run - T)).reshape(B * T_run, -1)
    rt = model._runtime(B, T_run, doc_start)
    with torch.enable_grad():
        res, delta = model._trunk(ids, rt)
        box = {"cot": cotangent.contiguous()}
        dummy = _HeadCotangentFn.apply(res, delta, model.head, rt, box)
    dummy.backward()


def make_ggn_matvec_fn_lm(model_def: Transformer, curvature_batch: Sequence, batch_stats=None):
    """optim/ggn_utils.py:69-153.  curvature_batch = (input_ids, labels, attn_mask): input_ids int [B, T] or
    [n_probe, B, T]; attn_mask None, the bool [.., B, T, T] block-diagonal causal mask of train_lm.py:107-131, or
    — cheaper — an int32 document-start tensor [.., B, T] (batch.doc_lengths_to_doc_start).

    Returns ggn_matvec(params, vec_pytree, rng_key=None) -> pytree of fp32 tensors with the reference's leaf names.
    `params` may be None (the module's current parameters) or a flax-named tree, which is loaded into the module."""
    if batch_stats is not None:
        raise NotImplementedError("batch_stats: the LM has no BatchNorm state (optim/ggn_utils.py:84)")
    if not isinstance(model_def, Transformer):
        raise TypeError("model_def must be a plaincv_b200 Transformer")
    model = model_def
    input_ids, labels, attn_mask = curvature_batch
    input_ids = torch.as_tensor(input_ids)
    if input_ids.dim() not in (2, 3):
        raise ValueError("input_ids must be [B, T] or [n_probe, B, T]")
    probes = [input_ids] if input_ids.dim() == 2 else [input_ids[i] for i in range(input_ids.shape[0])]
    masks = [None] * len(probes)
    if attn_mask is not None:
        am = torch.as_tensor(attn_mask)
        masks = [am] if input_ids.dim() == 2 else [am[i] for i in range(len(probes))]
    tangent = TangentBuffers(model)

    def _mask_kwargs(mk):
        if mk is None:
            return {}
        return {"attn_mask": mk} if mk.dtype == torch.bool else {"doc_start": mk}

    def ggn_matvec(params, vec_pytree, rng_key=None) -> VecTree:
        del rng_key
        if params is not None:
            model.load_flax_params(params)
        tangent.load(vec_pytree)
        model.flat_grad.zero_()
        n_probe = len(probes)
        for ids_b, mk in zip(probes, masks):
            # H (J vec) / (B*T) (:116-127); the mean over probe batches (:149-150) rides along in the same scale
            ggn_product_accumulate(model, ids_b, tangent, 1.0 / (ids_b.numel() * n_probe), **_mask_kwargs(mk))
        return model.export_flax_grads()

    return ggn_matvec
