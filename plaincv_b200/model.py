"""The models/LM causal Transformer (reference: models/LM/transformer.py, embedding.py, constructor.py) as a
torch.nn.Module whose every layer runs on the sm_100a kernels of libplaincv_b200.so.

Boundary honoured:
  * construct_model(cfg) -> (module, model_cfg): same cfg keys as models/LM/constructor.py:74-97;
  * parameter tree: module.state_dict() keys are the Flax paths with '/' -> '.', same shapes, kernels in flax's
    (in, out) layout, fp32 masters (param_dtype float32), no biases:
        embed_tokens.embedding [V,d]; layers_{i}.attn_norm.RMSNorm_0.scale [d]; layers_{i}.attn.w_qkv.kernel [d,3d];
        layers_{i}.attn.w_out.kernel [d,d]; layers_{i}.mlp_norm.RMSNorm_0.scale [d];
        layers_{i}.mlp.{fc_gate,fc_up}.kernel [d,hid]; layers_{i}.mlp.fc2.kernel [hid,d];
        out_norm.RMSNorm_0.scale [d]; lm_head.kernel [d,V]
  * module(input_ids, attn_mask=None, *, doc_lengths=None) -> logits [B,T,V]   (Transformer.__call__, :346-407)
  * module.loss_and_acc(input_ids, labels, doc_lengths=None) -> (loss, acc)    (fused head + _loss_and_acc,
    train_lm.py:181-186), never materialising fp32 logits.

Memory layout in HBM (B200-first, not a translation):
  * ONE flat fp32 master buffer holds every parameter in a padded layout (hid 2730 -> 2752, vocab 50257 -> 50304:
    128-byte row strides, whole L2 lines per TMA box row; fc_gate|fc_up share one [d, 2*hid_pad] matrix so gate and
    up are a single GEMM).
    The nn.Parameters are (possibly strided) views into it, so the pad lanes are never touched by an optimizer.
  * ONE flat bf16 buffer with the same offsets holds the compute copies; refreshing it after an optimizer step is a
    single cast kernel (this replaces flax's per-call promote_dtype of every kernel).
  * ONE flat fp32 gradient buffer with the same offsets; wgrad GEMMs red.add straight into it and param.grad are
    views of it, so gradient accumulation, zeroing and the data-parallel all-reduce are flat-buffer operations.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from fractions import Fraction
from typing import Callable, Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
from torch import nn

from . import ops
from .batch import doc_lengths_to_doc_start, doc_start_from_mask


@dataclass
class ModelConfig:
    """models/LM/transformer.py:13-26."""

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
    dtype: str = "bfloat16"
    param_dtype: str = "float32"

    @property
    def hidden_dim(self) -> int:
        return int(self.expand * self.dim)  # transformer.py:296

    @property
    def head_dim(self) -> int:
        return self.dim // self.n_heads


def _resolve_dtype(dtype_cfg, default: str = "float32") -> str:
    """models/LM/constructor.py:39-55."""
    if dtype_cfg is None:
        return default
    if isinstance(dtype_cfg, torch.dtype):
        dtype_cfg = {torch.float32: "float32", torch.float16: "float16", torch.bfloat16: "bfloat16"}.get(dtype_cfg)
    if isinstance(dtype_cfg, str):
        key = dtype_cfg.strip().lower()
        if key in ("float32", "float16", "bfloat16"):
            return key
    raise ValueError(f"Unsupported dtype '{dtype_cfg}'. Expected one of: float32, float16, bfloat16.")


def _cfg_get(cfg, key, default=None):
    if isinstance(cfg, dict):
        return cfg.get(key, default)
    return getattr(cfg, key, default)


def model_config_from_cfg(cfg) -> ModelConfig:
    """models/LM/constructor.py:74-97 (transformer branch)."""
    model = _cfg_get(cfg, "model")
    if model != "transformer":
        raise NotImplementedError(f"Not implemented model: {model}.")
    return ModelConfig(
        vocab_size=int(_cfg_get(cfg, "vocab_size")),
        dim=int(_cfg_get(cfg, "d_model")),
        expand=float(Fraction(str(_cfg_get(cfg, "expand")))),
        n_layers=int(_cfg_get(cfg, "n_layers")),
        n_heads=int(_cfg_get(cfg, "n_heads")),
        rmsnorm_eps=1e-6,
        mlp=_cfg_get(cfg, "mlp_class"),
        seq_len=int(_cfg_get(cfg, "seq_len")),
        tie_embeddings=bool(_cfg_get(cfg, "tie_embeddings", False)),
        rope_theta=float(_cfg_get(cfg, "rope_theta", 500000.0)),
        dtype=_resolve_dtype(_cfg_get(cfg, "dtype", "float32")),
        param_dtype=_resolve_dtype(_cfg_get(cfg, "param_dtype", "float32")),
    )


def _pad64(n: int) -> int:
    """Activation row strides are kept at whole 128-byte lines (64 bf16): a TMA box row that straddles two lines costs
    two L2 transactions, which the fused MLP epilogues feel (hid 2736 -> 2752: gate|up+SwiGLU GEMM 156 -> 149 us)."""
    return (n + 63) // 64 * 64


# ------------------------------------------------------------------------------------------------ flat layout
@dataclass
class _Slot:
    """A padded matrix [rows, ld] inside the flat buffers."""

    offset: int
    rows: int
    ld: int

    @property
    def numel(self) -> int:
        return self.rows * self.ld


@dataclass
class _ParamView:
    name: str  # flax path, '/'-joined
    slot: str
    row0: int
    col0: int
    shape: Tuple[int, ...]


class ParamLayout:
    """Where every reference parameter lives inside the flat master / bf16 / gradient buffers."""

    def __init__(self, cfg: ModelConfig):
        if cfg.dim % cfg.n_heads != 0:
            raise ValueError("dim must be divisible by n_heads")
        if cfg.mlp not in ("glu", "mlp", "mlp_relu_sq"):
            raise ValueError(f"Unknown mlp type: {cfg.mlp}")  # transformer.py:332
        self.gated = cfg.mlp == "glu"
        self.tied = bool(cfg.tie_embeddings)
        d, hid, V = cfg.dim, cfg.hidden_dim, cfg.vocab_size
        self.d, self.hid, self.V = d, hid, V
        self.hid_pad, self.V_pad = _pad64(hid), _pad64(V)
        if d % 8 != 0:
            raise ValueError("d_model must be a multiple of 8")
        self.slots: Dict[str, _Slot] = {}
        self.views: List[_ParamView] = []
        self.layer_ranges: List[Tuple[int, int]] = []  # [begin, end) element range of each layer's slots
        off = 0

        def add(slot_name: str, rows: int, ld: int) -> None:
            nonlocal off
            self.slots[slot_name] = _Slot(off, rows, ld)
            off += (rows * ld + 127) // 128 * 128

        # tied embeddings: the table doubles as the LM-head operand, so it gets the head's padded row count (pad rows
        # stay zero: zero logits in the pad columns, exactly as with the padded lm_head kernel)
        add("embed", self.V_pad if self.tied else V, d)
        self.views.append(_ParamView("embed_tokens/embedding", "embed", 0, 0, (V, d)))
        self.embed_range = (0, off)
        for i in range(cfg.n_layers):
            p = f"layers_{i}"
            begin = off
            add(f"{p}.attn_norm", 1, d)
            add(f"{p}.w_qkv", d, 3 * d)
            add(f"{p}.w_out", d, d)
            add(f"{p}.mlp_norm", 1, d)
            # glu: fc_gate | fc_up side by side (one GEMM); mlp / mlp_relu_sq: the slot holds fc1 alone
            add(f"{p}.w_gu", d, (2 if self.gated else 1) * self.hid_pad)
            add(f"{p}.w_down", self.hid_pad, d)
            self.layer_ranges.append((begin, off))
            self.views += [
                _ParamView(f"{p}/attn_norm/RMSNorm_0/scale", f"{p}.attn_norm", 0, 0, (d,)),
                _ParamView(f"{p}/attn/w_qkv/kernel", f"{p}.w_qkv", 0, 0, (d, 3 * d)),
                _ParamView(f"{p}/attn/w_out/kernel", f"{p}.w_out", 0, 0, (d, d)),
                _ParamView(f"{p}/mlp_norm/RMSNorm_0/scale", f"{p}.mlp_norm", 0, 0, (d,)),
            ]
            if self.gated:
                self.views += [_ParamView(f"{p}/mlp/fc_gate/kernel", f"{p}.w_gu", 0, 0, (d, hid)),
                               _ParamView(f"{p}/mlp/fc_up/kernel", f"{p}.w_gu", 0, self.hid_pad, (d, hid))]
            else:
                self.views.append(_ParamView(f"{p}/mlp/fc1/kernel", f"{p}.w_gu", 0, 0, (d, hid)))
            self.views.append(_ParamView(f"{p}/mlp/fc2/kernel", f"{p}.w_down", 0, 0, (hid, d)))
        begin = off
        add("out_norm", 1, d)
        self.views.append(_ParamView("out_norm/RMSNorm_0/scale", "out_norm", 0, 0, (d,)))
        if not self.tied:  # tie_embeddings: logits = embed.attend(x) (transformer.py:394-396), no lm_head parameter
            add("w_lm", d, self.V_pad)
            self.views.append(_ParamView("lm_head/kernel", "w_lm", 0, 0, (d, V)))
        self.head_range = (begin, off)
        self.total = off

    def slot_view(self, flat: torch.Tensor, slot_name: str) -> torch.Tensor:
        s = self.slots[slot_name]
        return flat[s.offset: s.offset + s.numel].view(s.rows, s.ld)

    def param_view(self, flat: torch.Tensor, v: _ParamView) -> torch.Tensor:
        m = self.slot_view(flat, v.slot)
        if len(v.shape) == 1:
            return m[0, : v.shape[0]]
        return m[v.row0: v.row0 + v.shape[0], v.col0: v.col0 + v.shape[1]]


class _Node(nn.Module):
    """Anonymous container so that state_dict keys mirror the Flax tree ('layers_0.attn.w_qkv.kernel')."""


class _LayerHandles:
    """Raw tensors one block's kernels read/write (no nn.Module machinery on the hot path)."""

    __slots__ = ("attn_norm", "mlp_norm", "w_qkv", "w_out", "w_gu", "w_down", "g_attn_norm", "g_mlp_norm", "g_qkv",
                 "g_out", "g_gu", "g_down", "mlp", "on_grads_ready")


class _HeadHandles:
    __slots__ = ("out_norm", "w_lm", "g_out_norm", "g_lm", "vocab", "v_pad", "tied", "on_grads_ready")


_T_TILE = 128  # query / key tile of the attention kernels


def _round_up(n: int, m: int) -> int:
    return (n + m - 1) // m * m


class _Runtime:
    """Per-call constants shared by every block of one forward pass."""

    __slots__ = ("eps", "cos_sin", "doc_start", "B", "T", "n_heads", "tape")


# ------------------------------------------------------------------------------------------------ the module
class Transformer(nn.Module):
    def __init__(self, cfg: ModelConfig, device: Optional[torch.device] = None, seed: int = 0):
        super().__init__()
        if _resolve_dtype(cfg.dtype) != "bfloat16":
            raise NotImplementedError("the B200 path computes in bfloat16 (cfg.dtype); fp32 is the oracle's job")
        if _resolve_dtype(cfg.param_dtype) != "float32":
            raise NotImplementedError("param_dtype must be float32 (fp32 masters + bf16 compute copies)")
        if cfg.head_dim != 64:
            raise NotImplementedError("attention kernels are specialised for head_dim = 64 (every reference config)")
        self.cfg = cfg
        self.layout = ParamLayout(cfg)
        device = torch.device(device if device is not None else "cuda")
        if device.type != "cuda":
            raise RuntimeError("plaincv_b200.Transformer needs a CUDA device: there is no CPU fallback "
                               "(the CPU restatement lives in oracle/ and is test infrastructure only)")
        L = self.layout
        self.flat_master = torch.zeros(L.total, dtype=torch.float32, device=device)
        self.flat_bf16 = torch.zeros(L.total, dtype=torch.bfloat16, device=device)
        self.flat_grad = torch.zeros(L.total, dtype=torch.float32, device=device)
        self._flax_names: Dict[str, str] = {}
        self._grad_views: Dict[str, torch.Tensor] = {}
        for v in L.views:
            p = nn.Parameter(L.param_view(self.flat_master, v), requires_grad=True)
            self._grad_views[v.name] = L.param_view(self.flat_grad, v)
            parts = v.name.split("/")
            node: nn.Module = self
            for part in parts[:-1]:
                if not hasattr(node, part):
                    node.add_module(part, _Node())
                node = getattr(node, part)
            node.register_parameter(parts[-1], p)
            self._flax_names[v.name] = ".".join(parts)
        self._params_by_flax = dict(zip(self._flax_names.keys(), [self.get_parameter(n) for n in self._flax_names.values()]))
        # (rows up to the next multiple of the attention tile: shorter or ragged sequences run padded, _prep_padded)
        self.cos_sin = ops.rope_table(cfg.head_dim, _round_up(cfg.seq_len, _T_TILE), cfg.rope_theta, device)
        self._build_handles()
        self.reset_parameters(seed)
        self._attach_grads()

    # -------------------------------------------------------------------------------- parameters
    def _build_handles(self) -> None:
        L = self.layout
        self.layers: List[_LayerHandles] = []
        for i in range(self.cfg.n_layers):
            p = f"layers_{i}"
            h = _LayerHandles()
            h.attn_norm = L.slot_view(self.flat_master, f"{p}.attn_norm").view(-1)
            h.mlp_norm = L.slot_view(self.flat_master, f"{p}.mlp_norm").view(-1)
            h.g_attn_norm = L.slot_view(self.flat_grad, f"{p}.attn_norm").view(-1)
            h.g_mlp_norm = L.slot_view(self.flat_grad, f"{p}.mlp_norm").view(-1)
            for short in ("qkv", "out", "gu", "down"):
                setattr(h, f"w_{short}", L.slot_view(self.flat_bf16, f"{p}.w_{short}"))
                setattr(h, f"g_{short}", L.slot_view(self.flat_grad, f"{p}.w_{short}"))
            h.mlp = self.cfg.mlp
            h.on_grads_ready = None
            self.layers.append(h)
        hd = _HeadHandles()
        hd.out_norm = L.slot_view(self.flat_master, "out_norm").view(-1)
        hd.g_out_norm = L.slot_view(self.flat_grad, "out_norm").view(-1)
        hd.tied = L.tied
        if L.tied:  # the LM head reads the embedding table [V, d] and adds its gradient to the table's
            hd.w_lm = L.slot_view(self.flat_bf16, "embed")
            hd.g_lm = L.slot_view(self.flat_grad, "embed")
        else:
            hd.w_lm = L.slot_view(self.flat_bf16, "w_lm")
            hd.g_lm = L.slot_view(self.flat_grad, "w_lm")
        hd.vocab = self.cfg.vocab_size
        hd.v_pad = L.V_pad
        hd.on_grads_ready = None
        self.head = hd
        self.embed_bf16 = L.slot_view(self.flat_bf16, "embed")
        self.embed_grad = L.slot_view(self.flat_grad, "embed")

    @torch.no_grad()
    def reset_parameters(self, seed: int = 0) -> None:
        """N(0, 0.02); w_out / fc2: N(0, 0.02/sqrt(2L)); norm scales 1 (transformer.py:185-187, :297-298, :359)."""
        g = torch.Generator().manual_seed(seed)
        resid_std = 0.02 / math.sqrt(2 * self.cfg.n_layers)
        for name, p in self._params_by_flax.items():
            if name.endswith("scale"):
                p.fill_(1.0)
            else:
                std = resid_std if ("w_out" in name or "fc2" in name) else 0.02
                p.copy_((torch.randn(tuple(p.shape), generator=g, dtype=torch.float32) * std).to(p.device))
        self.sync_compute_weights()

    @torch.no_grad()
    def load_flax_params(self, params: Dict[str, torch.Tensor]) -> None:
        """Load a {'layers_0/attn/w_qkv/kernel': tensor, ...} tree (the reference's variables['params'] flattened)."""
        missing = set(self._params_by_flax) - set(params)
        extra = set(params) - set(self._params_by_flax)
        if missing or extra:
            raise KeyError(f"parameter tree mismatch: missing={sorted(missing)[:4]} unexpected={sorted(extra)[:4]}")
        for name, p in self._params_by_flax.items():
            src = torch.as_tensor(params[name])
            if tuple(src.shape) != tuple(p.shape):
                raise ValueError(f"{name}: expected shape {tuple(p.shape)}, got {tuple(src.shape)}")
            p.copy_(src.to(device=p.device, dtype=torch.float32))
        self.sync_compute_weights()

    def export_flax_params(self) -> Dict[str, torch.Tensor]:
        return {k: p.detach().clone() for k, p in self._params_by_flax.items()}

    def export_flax_grads(self) -> Dict[str, torch.Tensor]:
        return {k: g.detach().clone() for k, g in self._grad_views.items()}

    def sync_compute_weights(self) -> None:
        """fp32 masters -> bf16 compute copies: one cast kernel over the flat buffer.  Call after every optimizer
        step (the engine does); the reference re-casts every kernel inside every forward instead."""
        n = self.layout.total
        ops.cast_pad_bf16(self.flat_master, self.flat_bf16, 1, n, n, n)

    def _attach_grads(self) -> None:
        for name, p in self._params_by_flax.items():
            if p.grad is None or p.grad.data_ptr() != self._grad_views[name].data_ptr():
                p.grad = self._grad_views[name]

    def zero_grad(self, set_to_none: bool = False) -> None:  # noqa: D401 - keep nn.Module's signature
        """Gradients live in the flat buffer; zeroing is one memset and .grad stays attached."""
        self.flat_grad.zero_()
        self._attach_grads()

    def count_params(self) -> Tuple[int, int]:
        """(all, non-embedding) — constructor.py:12-36."""
        total = sum(int(p.numel()) for p in self._params_by_flax.values())
        return total, total - int(self._params_by_flax["embed_tokens/embedding"].numel())

    # -------------------------------------------------------------------------------- forward
    def _runtime(self, B: int, T: int, doc_start: Optional[torch.Tensor]) -> _Runtime:
        rt = _Runtime()
        rt.eps, rt.cos_sin, rt.doc_start = self.cfg.rmsnorm_eps, self.cos_sin, doc_start
        rt.B, rt.T, rt.n_heads = B, T, self.cfg.n_heads
        rt.tape = None  # a list: every block appends its saved activations (read by the tangent pass, curvature.py)
        return rt

    def _prep_inputs(self, input_ids, attn_mask, doc_lengths, doc_start):
        if input_ids.dim() != 2:
            raise ValueError("input_ids must be [B, T]")
        B, T = input_ids.shape
        if T > self.cfg.seq_len:
            raise ValueError(f"sequence length {T} exceeds cfg.seq_len={self.cfg.seq_len}")
        dev = self.flat_master.device
        ids = input_ids.to(device=dev, dtype=torch.int32, non_blocking=True).contiguous().view(-1)
        if doc_start is None and doc_lengths is not None:
            doc_start = torch.from_numpy(doc_lengths_to_doc_start(doc_lengths, T))
        if doc_start is None and attn_mask is not None:
            doc_start = doc_start_from_mask(attn_mask)
        if doc_start is not None:
            doc_start = doc_start.to(device=dev, dtype=torch.int32, non_blocking=True).contiguous().view(-1)
            if doc_start.numel() != B * T:
                raise ValueError("doc_start must have B*T entries")
        return ids, B, T, doc_start

    def _prep_padded(self, input_ids, attn_mask, doc_lengths, doc_start):
        """_prep_inputs for any T <= seq_len (the reference slices freqs_cis to the batch's length,
        transformer.py:373-374): the attention kernels tile queries and keys by 128, so a ragged T runs as
        T_run = T rounded up, every sequence right-padded with token 0.  Causality keeps real positions from ever seeing
        a pad key, position t keeps rotation t (rows are b * T_run + t), so real rows are exactly what the un-padded
        computation gives; the callers drop the pad rows (logits) or keep them out of the loss and zero their
        d(logits) (ops._ce_means).  -> (ids, B, T, T_run, doc_start) with ids / doc_start of B * T_run entries."""
        ids, B, T, doc_start = self._prep_inputs(input_ids, attn_mask, doc_lengths, doc_start)
        T_run = _round_up(T, _T_TILE)
        if T_run != T:
            ids = torch.nn.functional.pad(ids.view(B, T), (0, T_run - T)).contiguous().view(-1)
            if doc_start is not None:  # every pad position is a document of its own
                tail = torch.arange(T, T_run, device=ids.device, dtype=torch.int32).expand(B, T_run - T)
                doc_start = torch.cat([doc_start.view(B, T), tail], dim=1).contiguous().view(-1)
        return ids, B, T, T_run, doc_start

    def _trunk(self, ids: torch.Tensor, rt: _Runtime):
        self._attach_grads()
        L = self.layout
        x = ops.EmbedFn.apply(self._params_by_flax["embed_tokens/embedding"], ids, self.embed_bf16, self.embed_grad,
                              L.d, L.d)
        res, delta = x, None
        for layer in self.layers:
            res, delta = ops.BlockFn.apply(res, delta, layer, rt)
        return res, delta

    def forward(self, input_ids: torch.Tensor, attn_mask: Optional[torch.Tensor] = None, *,
                doc_lengths: Optional[Sequence[Sequence[int]]] = None, doc_start: Optional[torch.Tensor] = None,
                deterministic: bool = True) -> torch.Tensor:
        """Transformer.__call__ (transformer.py:346-407) -> logits [B, T, V] bf16.  `attn_mask`, if given, must be the
        bool [B,T,T] block-diagonal causal mask train_lm.py builds; it is converted to document starts."""
        ids, B, T, T_run, doc_start = self._prep_padded(input_ids, attn_mask, doc_lengths, doc_start)
        rt = self._runtime(B, T_run, doc_start)
        res, delta = self._trunk(ids, rt)
        res_f, hf = ops.RMSNormResidualFn.apply(res, delta, self.head.out_norm, self.head.g_out_norm, rt.eps)
        if self.head.tied:
            if torch.is_grad_enabled() and res.requires_grad:
                raise NotImplementedError("tie_embeddings: differentiate through loss_and_acc (the fused head), "
                                          "module(input_ids) returns logits for inference only")
            logits = ops.head_logits(hf, self.head)
        else:
            logits = ops.LinearFn.apply(hf, self.head.w_lm, self.head.g_lm)
        return logits.view(B, T_run, self.layout.V_pad)[:, :T, : self.cfg.vocab_size]

    def loss_and_acc(self, input_ids: torch.Tensor, labels: torch.Tensor, *,
                     doc_lengths: Optional[Sequence[Sequence[int]]] = None, doc_start: Optional[torch.Tensor] = None,
                     attn_mask: Optional[torch.Tensor] = None, loss_scale: float = 1.0, unit_grad: bool = False):
        """model.apply + _loss_and_acc (train_lm.py:259-265, :181-186) -> (loss, acc) 0-dim fp32 tensors.
        `loss_scale` pre-scales the gradients (1/grad_accumulation_steps); `unit_grad=True` promises that
        loss.backward() is called with the default unit gradient (saves a device->host read of it)."""
        ids, B, T, T_run, doc_start = self._prep_padded(input_ids, attn_mask, doc_lengths, doc_start)
        if tuple(labels.shape) != (B, T):
            raise ValueError("labels must match input_ids' shape")
        lab = labels.to(device=ids.device, dtype=torch.int32, non_blocking=True).contiguous()
        pad = None
        if T_run != T:
            lab = torch.nn.functional.pad(lab, (0, T_run - T)).contiguous()
            pad = (B, T, T_run)
        rt = self._runtime(B, T_run, doc_start)
        res, delta = self._trunk(ids, rt)
        out2 = ops.HeadLossFn.apply(res, delta, lab.view(-1), self.head, rt, loss_scale, unit_grad, pad)
        return out2[0], out2[1]


def construct_model(cfg, device: Optional[torch.device] = None, seed: Optional[int] = None):
    """models/LM/constructor.py:58-104 -> (module, model_cfg).  The reference also returns `variables`; here the
    parameters live in the module (module.export_flax_params() gives the same tree)."""
    model_cfg = model_config_from_cfg(cfg)
    if seed is None:
        seed = int(_cfg_get(cfg, "seed", 0) or 0)
    module = Transformer(model_cfg, device=device, seed=seed)
    return module, model_cfg
