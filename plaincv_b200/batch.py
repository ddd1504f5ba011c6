"""Host-side batch preparation: the contract of train_lm.py:_prepare_batch (:134-148), _build_attn_mask (:107-131),
_trim_last_token (:97-104) and _reshape_for_pmap (:151-159) — with one B200-first change: the intra-document mask is
carried as `doc_start[b, t]` (first position of the document containing token t) instead of a bool [B, T, T] tensor
(33.5 MB per device-step at T=2048, built by a Python loop + torch.block_diag in the reference).  The attention kernels
consume doc_start directly; `doc_start_from_mask` accepts the reference's dense mask for API parity.

Pure numpy / torch-CPU: covered by the `not gpu` tests.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch


def trim_last_token(doc_boundaries: Sequence[int]) -> List[int]:
    """train_lm.py:97-104: inputs drop the final token, so the last document shrinks by one (and vanishes at 0)."""
    trimmed = list(doc_boundaries)
    if not trimmed:
        return trimmed
    trimmed[-1] -= 1
    if trimmed[-1] <= 0:
        trimmed.pop()
    return trimmed


def doc_lengths_to_doc_start(docs_lengths: Sequence[Sequence[int]], seq_len: int) -> np.ndarray:
    """docs_lengths (per row, summing to seq_len+1) -> int32 [B, seq_len] document-start positions.
    Equivalent to intra_doc_causal_mask(_trim_last_token(lengths), seq_len) (data_prep_utils.py:14-43):
    mask[t, s] = (doc_start[t] <= s <= t)."""
    out = np.empty((len(docs_lengths), seq_len), dtype=np.int32)
    for b, boundaries in enumerate(docs_lengths):
        boundaries = trim_last_token([int(x) for x in boundaries])
        if sum(boundaries) != seq_len:
            raise ValueError(f"Sum(doc_boundaries)={sum(boundaries)} != seq_len={seq_len}.")
        if any(n <= 0 for n in boundaries):
            raise ValueError("document lengths must be positive")
        starts = np.cumsum([0] + boundaries[:-1])
        out[b] = np.repeat(starts, boundaries)
    return out


def doc_start_from_mask(attn_mask: torch.Tensor) -> torch.Tensor:
    """bool [B, T, T] block-diagonal causal mask (what train_lm.py:_build_attn_mask produces) -> int32 [B, T].
    Raises ValueError for masks that are not of that family (the kernels implement causal ∩ document blocks only;
    additive-bias masks, transformer.py:227-229, are unused by train_lm.py)."""
    if attn_mask.dtype != torch.bool or attn_mask.dim() != 3 or attn_mask.shape[1] != attn_mask.shape[2]:
        raise ValueError("attn_mask must be a bool [B, T, T] tensor")
    B, T, _ = attn_mask.shape
    first = torch.argmax(attn_mask.to(torch.uint8), dim=-1).to(torch.int32)  # first True per row
    s = torch.arange(T, device=attn_mask.device)
    rebuilt = (s[None, None, :] >= first[:, :, None]) & (s[None, None, :] <= s[None, :, None])
    if not torch.equal(rebuilt, attn_mask):
        raise ValueError("attn_mask is not a block-diagonal causal (intra-document) mask")
    if T > 1 and bool((first[:, 1:] < first[:, :-1]).any()):
        raise ValueError("attn_mask document starts must be non-decreasing")
    return first


def dense_mask_from_doc_start(doc_start: np.ndarray) -> np.ndarray:
    """Inverse of the above, for tests: bool [B, T, T]."""
    ds = np.asarray(doc_start)
    T = ds.shape[1]
    s = np.arange(T)
    return (s[None, None, :] >= ds[:, :, None]) & (s[None, None, :] <= s[None, :, None])


def check_token_range(input_ids: torch.Tensor, vocab_size: int) -> None:
    """The kernels use token ids as raw row indices (embedding gather, its scatter-add transpose, the label pick of
    the loss) — an id outside [0, vocab_size), e.g. data tokenised with another vocabulary, would read or red.add
    outside the table.  XLA clamps gathers silently; here the batch is rejected on the host (one min/max pass)."""
    if input_ids.numel() == 0:
        return
    lo, hi = int(input_ids.min()), int(input_ids.max())
    if lo < 0 or hi >= vocab_size:
        raise ValueError(f"token ids must lie in [0, {vocab_size}); batch holds [{lo}, {hi}]")


def prepare_batch(batch: dict, seq_len: int, use_doc_mask: bool,
                  vocab_size: Optional[int] = None) -> Tuple[torch.Tensor, torch.Tensor, Optional[torch.Tensor]]:
    """train_lm.py:134-148 -> (inputs int32 [B,T], labels int32 [B,T], doc_start int32 [B,T] | None), host tensors.
    vocab_size (optional): reject token ids outside [0, vocab_size) (check_token_range)."""
    input_ids = torch.as_tensor(np.asarray(batch["input_ids"])).to(torch.int32)
    if input_ids.dim() != 2 or input_ids.shape[1] != seq_len + 1:
        got = input_ids.shape[1] if input_ids.dim() == 2 else tuple(input_ids.shape)
        raise ValueError(f"Expected input_ids length {seq_len + 1} (seq_len+1) but got {got}.")
    if vocab_size is not None:
        check_token_range(input_ids, vocab_size)
    inputs = input_ids[:, :-1]
    labels = input_ids[:, 1:]
    doc_start = None
    if use_doc_mask:
        docs_lengths = batch.get("docs_lengths", None)
        if docs_lengths is None:
            raise ValueError("intra_doc_masking=True but docs_lengths not found in batch.")
        if len(docs_lengths) != input_ids.shape[0]:
            raise ValueError("docs_lengths must have one entry per sequence")
        doc_start = torch.from_numpy(doc_lengths_to_doc_start(docs_lengths, seq_len))
    return inputs, labels, doc_start


def shard_for_rank(x: Optional[torch.Tensor], rank: int, world_size: int) -> Optional[torch.Tensor]:
    """The per-device slice of _reshape_for_pmap (train_lm.py:151-159): contiguous split of the leading dimension."""
    if x is None:
        return None
    if x.shape[0] % world_size != 0:
        raise ValueError(f"Leading batch dimension {x.shape[0]} is not divisible by n_devices={world_size}.")
    per = x.shape[0] // world_size
    return x[rank * per: (rank + 1) * per]


def synthetic_batch(batch_size: int, seq_len: int, vocab_size: int, seed: int = 1234,
                    doc_mean: Optional[int] = None) -> dict:
    """A synthetic batch in the loader's contract (data/lm_loader.py:32-38): `input_ids` int32 [B, T+1] uniform in
    [0, V) and, when doc_mean is given, `docs_lengths` (list per row, summing to T+1, exponential lengths — the shape
    of concat_chunck's output, data_prep_utils.py:117-141).  The reference has no synthetic-data mode (README TODO);
    bench.py and the tests feed this straight into the _prepare_batch contract."""
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
