"""Batch feeding for the hot path: what sits between the token files on disk and
`Engine.compute_grads` in the reference —

  * `data/lm_loader.py:17-85`      datasets.load_from_disk + per-process contiguous shard + torch DataLoader (collate:
                                   `input_ids` stacked to [B, T+1], `docs_lengths` kept as lists, drop_last=True)
  * `data/lm_datasampler.py:19-160` the index streams (sequential / seeded random, resumable from a micro-step count;
                                   per-rank contiguous blocks with an epoch-seeded shuffle)
  * `train_lm.py:356-399`          `_next_batch`: wrap around at the end of the data, merge grouped batches
  * `train_lm.py:134-148, 107-131` `_prepare_batch` / `_build_attn_mask` (here: `batch.prepare_batch`, document starts)

— restated so that a > 300 k tokens/s step never waits for its input: rows come straight out of the Arrow files that
`datasets.Dataset.save_to_disk` wrote (memory-mapped, no `datasets` import, no worker processes), a background thread
assembles micro-batches into a ring of pinned host buffers and queues their host-to-device copies on a side stream, and
the training loop receives device tensors fenced by an event.  On a machine without CUDA the feeder yields the same
tensors on the host (that is how the CPU tests drive it).
"""
from __future__ import annotations

import glob
import json
import os
import queue
import threading
from typing import Dict, Iterator, List, Optional, Sequence

import numpy as np
import torch

from .batch import prepare_batch


# ------------------------------------------------------------------------------------------------ dataset on disk
class ArrowTokenDataset:
    """Rows of a `datasets.Dataset.save_to_disk` directory: `state.json` lists the Arrow stream files, each holding
    the columns `input_ids` (list<int>, T+1 tokens per row) and optionally `docs_lengths` (list<int>, summing to T+1).
    Mirrors the part of `datasets.Dataset` the loader uses: len, row access, `column_names`, `shard(contiguous=True)`,
    `take`."""

    def __init__(self, path: str, _table=None, _start: int = 0, _stop: Optional[int] = None):
        import pyarrow as pa
        if _table is None:
            state_file = os.path.join(path, "state.json")
            if not os.path.isfile(state_file):
                raise ValueError("dataset should be a datasets.Dataset")  # lm_loader.py:19-20 (not a saved Dataset)
            with open(state_file) as f:
                state = json.load(f)
            files = [os.path.join(path, d["filename"]) for d in state.get("_data_files", [])]
            if not files:
                files = sorted(glob.glob(os.path.join(path, "*.arrow")))
            tables = []
            for fn in files:
                src = pa.memory_map(fn, "r")
                try:
                    tables.append(pa.ipc.open_stream(src).read_all())
                except pa.ArrowInvalid:
                    tables.append(pa.ipc.open_file(pa.memory_map(fn, "r")).read_all())
            _table = pa.concat_tables(tables) if len(tables) > 1 else tables[0]
        self.path = path
        self._table = _table
        self._start = _start
        self._stop = _table.num_rows if _stop is None else _stop
        self.column_names: List[str] = list(_table.column_names)
        if "input_ids" not in self.column_names:
            raise ValueError("dataset has no 'input_ids' column")
        self._ids2d = None  # [rows, T+1] view when every row has the same length (the reference's datasets do)
        col = _table.column("input_ids").combine_chunks()
        if isinstance(col, pa.ChunkedArray):
            col = col.chunk(0) if col.num_chunks == 1 else pa.concat_arrays(col.chunks)
        self._ids_col = col
        if _table.num_rows > 0:
            offs = col.offsets.to_numpy()
            width = int(offs[1] - offs[0])
            if np.all(np.diff(offs) == width) and col.values.null_count == 0:
                flat = col.values.to_numpy(zero_copy_only=False)
                self._ids2d = flat[int(offs[0]): int(offs[0]) + width * _table.num_rows].reshape(_table.num_rows, width)
        self._doc_col = None
        if "docs_lengths" in self.column_names:
            dcol = _table.column("docs_lengths").combine_chunks()
            if isinstance(dcol, pa.ChunkedArray):
                dcol = dcol.chunk(0) if dcol.num_chunks == 1 else pa.concat_arrays(dcol.chunks)
            self._doc_offs = dcol.offsets.to_numpy()
            self._doc_vals = dcol.values.to_numpy(zero_copy_only=False)
            self._doc_col = dcol

    def __len__(self) -> int:
        return self._stop - self._start

    def _row_ids(self, i: int) -> np.ndarray:
        if self._ids2d is not None:
            return self._ids2d[i]
        return np.asarray(self._ids_col[i].as_py())

    def __getitem__(self, i: int) -> Dict[str, object]:
        if i < 0:
            i += len(self)
        if not 0 <= i < len(self):
            raise IndexError(i)
        j = self._start + i
        out: Dict[str, object] = {"input_ids": self._row_ids(j)}
        if self._doc_col is not None:
            out["docs_lengths"] = self._doc_vals[self._doc_offs[j]: self._doc_offs[j + 1]].tolist()
        return out

    def rows(self, idx: Sequence[int]) -> Dict[str, object]:
        """The collate of lm_loader.py:30-37 for a list of row indices: `input_ids` stacked, `docs_lengths` as lists."""
        j = np.asarray(idx, dtype=np.int64) + self._start
        if self._ids2d is not None:
            ids = self._ids2d[j]
        else:
            ids = np.stack([self._row_ids(int(k)) for k in j], axis=0)
        out: Dict[str, object] = {"input_ids": ids}
        if self._doc_col is not None:
            out["docs_lengths"] = [self._doc_vals[self._doc_offs[k]: self._doc_offs[k + 1]].tolist() for k in j]
        return out

    def shard(self, num_shards: int, index: int, contiguous: bool = True) -> "ArrowTokenDataset":
        """datasets.Dataset.shard(contiguous=True): the first n % num_shards shards get one extra row
        (lm_loader.py:23-28 — one shard per process)."""
        if not contiguous:
            raise NotImplementedError("only contiguous shards (what the reference's loader asks for)")
        if not 0 <= index < num_shards:
            raise ValueError("index should be in [0, num_shards-1]")
        n = len(self)
        div, mod = divmod(n, num_shards)
        start = div * index + min(index, mod)
        stop = start + div + (1 if index < mod else 0)
        return ArrowTokenDataset(self.path, self._table, self._start + start, self._start + stop)

    def take(self, n: int) -> "ArrowTokenDataset":
        return ArrowTokenDataset(self.path, self._table, self._start, self._start + min(int(n), len(self)))


# ------------------------------------------------------------------------------------------------ index streams
class InMemoryTokenDataset:
    """Rows held in host memory behind the same interface as ArrowTokenDataset (`len`, `[i]`, `rows(idx)`): the
    synthetic-data mode the reference's README lists as a TODO.  `input_ids` int [n, T+1]; `docs_lengths` optional
    list (one list of document lengths per row, each summing to T+1 — lm_loader.py:32-38)."""

    def __init__(self, input_ids, docs_lengths: Optional[Sequence[Sequence[int]]] = None):
        self._ids2d = np.ascontiguousarray(np.asarray(input_ids))
        if self._ids2d.ndim != 2:
            raise ValueError("input_ids must be [n_rows, seq_len + 1]")
        if docs_lengths is not None and len(docs_lengths) != len(self._ids2d):
            raise ValueError("docs_lengths must have one entry per row")
        self._docs = None if docs_lengths is None else [list(map(int, d)) for d in docs_lengths]

    @classmethod
    def synthetic(cls, n_rows: int, seq_len: int, vocab_size: int, seed: int = 1234,
                  doc_mean: Optional[int] = None) -> "InMemoryTokenDataset":
        from .batch import synthetic_batch
        b = synthetic_batch(n_rows, seq_len, vocab_size, seed=seed, doc_mean=doc_mean)
        return cls(b["input_ids"], b.get("docs_lengths"))

    def __len__(self) -> int:
        return len(self._ids2d)

    def __getitem__(self, i: int) -> Dict[str, object]:
        if i < 0:
            i += len(self)
        if not 0 <= i < len(self):
            raise IndexError(i)
        out: Dict[str, object] = {"input_ids": self._ids2d[i]}
        if self._docs is not None:
            out["docs_lengths"] = list(self._docs[i])
        return out

    def rows(self, idx: Sequence[int]) -> Dict[str, object]:
        j = np.asarray(idx, dtype=np.int64)
        out: Dict[str, object] = {"input_ids": self._ids2d[j]}
        if self._docs is not None:
            out["docs_lengths"] = [list(self._docs[int(k)]) for k in j]
        return out


def epoch_indices(n: int, sampler: str = "sequential", *, batch_size: int = 1, seed: Optional[int] = None,
                  start_idx: int = 0, rng: Optional[np.random.Generator] = None) -> np.ndarray:
    """One epoch's row order for the samplers `_get_sampler_jax` offers (lm_loader.py:88-118):
    'sequential' / 'stateful_sequential' (lm_datasampler.py:19-31): range(start_idx * batch_size, n);
    'stateful_random' (lm_datasampler.py:34-70): rng.permutation(n)[start_idx * batch_size:] with a *persistent*
    numpy Generator (pass the same `rng` for every epoch; it is seeded once with `seed`).  'random' (torch
    RandomSampler) draws from torch's generator and is not resumable; it maps to a seeded numpy permutation here."""
    start = int(start_idx) * int(batch_size)
    if sampler in ("sequential", "stateful_sequential"):
        return np.arange(start, n, dtype=np.int64)
    if sampler in ("stateful_random", "random"):
        if rng is None:
            if seed is None:
                if sampler == "stateful_random":
                    raise ValueError("Seed must be set if shuffle=True in a stateful sampler.")
                seed = 0
            rng = np.random.default_rng(int(seed))
        return rng.permutation(np.arange(n, dtype=np.int64))[start:]
    raise NotImplementedError(f"Sampler {sampler} is not implemented.")


def distributed_indices(n: int, rank: int, num_replicas: int, batch_size: int, *, seed: int = 0, epoch: int = 0,
                        start_iter: int = 0, shuffle_within_rank: bool = True, drop_last: bool = True) -> np.ndarray:
    """StatefulJaxDistributedSampler.__iter__ (lm_datasampler.py:73-146): rank r owns the contiguous block
    [r * num_samples, (r+1) * num_samples), shuffled within the rank by default_rng(seed + epoch), resumed at
    start_iter * batch_size."""
    if batch_size <= 0:
        raise ValueError("batch_size must be > 0")
    total = n - (n % num_replicas) if drop_last else int(-(-n // num_replicas) * num_replicas)
    num_samples = total // num_replicas
    idx = np.arange(rank * num_samples, (rank + 1) * num_samples, dtype=np.int64)
    if not drop_last:
        idx = idx % n
    if shuffle_within_rank:
        idx = idx[np.random.default_rng(int(seed) + int(epoch)).permutation(num_samples)]
    return idx[int(start_iter) * int(batch_size):]


# ------------------------------------------------------------------------------------------------ the feeder
class DeviceBatch:
    """One micro-batch staged for `Engine.compute_grads`: int32 inputs / labels [B, T] (and doc_start [B, T] when the
    run is document-masked) on the target device, plus the event that fences their copies."""

    __slots__ = ("inputs", "labels", "doc_start", "ready", "_slot")

    def __init__(self, inputs, labels, doc_start, ready, slot):
        self.inputs, self.labels, self.doc_start, self.ready, self._slot = inputs, labels, doc_start, ready, slot


class BatchFeeder:
    """Endless iterator over micro-batches of `dataset` (`_next_batch`, train_lm.py:370-380: wrap around at the end;
    `drop_last=True`, lm_loader.py:45), prepared (`_prepare_batch`: inputs = ids[:, :-1], labels = ids[:, 1:], document
    starts instead of the [B,T,T] mask) and staged `depth` batches ahead of the consumer.

    state(): (epoch, batches consumed in this epoch) — what `resume_step * grad_accumulation_steps` encodes in the
    reference's stateful samplers; pass it back as `start_idx` (and the same seed) to continue the stream."""

    def __init__(self, dataset: ArrowTokenDataset, micro_batch_size: int, seq_len: int, *, use_doc_mask: bool = False,
                 sampler: str = "sequential", sampler_seed: Optional[int] = None, start_idx: int = 0,
                 device: Optional[torch.device] = None, depth: int = 3, reapply_start_idx: bool = True,
                 vocab_size: Optional[int] = None):
        """reapply_start_idx=True mirrors the reference's stateful samplers, whose `__iter__` skips the first
        `start_idx * batch_size` samples on EVERY epoch after a resume, not only the first
        (lm_datasampler.py:24-27, 59-63); False applies the offset to the first epoch only.
        vocab_size, when given, makes the producer reject rows with token ids outside [0, vocab_size) — the kernels
        index the embedding table and the logits with them unchecked."""
        if micro_batch_size <= 0 or len(dataset) < micro_batch_size:
            raise ValueError("dataset holds fewer rows than one micro-batch")
        self.ds, self.B, self.T = dataset, int(micro_batch_size), int(seq_len)
        self.use_doc_mask = bool(use_doc_mask)
        self.sampler, self.seed, self.start_idx = sampler, sampler_seed, int(start_idx)
        self.device = torch.device(device) if device is not None else None
        self.cuda = self.device is not None and self.device.type == "cuda"
        if self.cuda and self.device.index is None:  # the worker thread needs an explicit ordinal
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.depth = max(1, int(depth))
        self.reapply_start_idx = bool(reapply_start_idx)
        self.vocab_size = None if vocab_size is None else int(vocab_size)
        if self.start_idx * self.B > len(dataset) - self.B:
            raise ValueError("start_idx leaves less than one micro-batch in an epoch")
        self._rng = np.random.default_rng(int(sampler_seed)) if sampler in ("stateful_random", "random") and \
            sampler_seed is not None else None
        self._q: "queue.Queue" = queue.Queue(maxsize=self.depth)
        self._free: "queue.Queue" = queue.Queue()
        self._stop = threading.Event()
        self._err: Optional[BaseException] = None
        self.epoch, self.consumed = 0, self.start_idx
        n_buf = self.depth + 2  # staged + the one the consumer holds + the one being filled
        self._slots = []
        self._slot_copy_done: List[Optional["torch.cuda.Event"]] = [None] * n_buf
        for k in range(n_buf):
            host = [torch.empty((self.B, self.T), dtype=torch.int32) for _ in range(3 if self.use_doc_mask else 2)]
            if self.cuda:
                host = [h.pin_memory() for h in host]
                dev = [torch.empty((self.B, self.T), dtype=torch.int32, device=self.device) for _ in host]
            else:
                dev = host
            self._slots.append((host, dev))
            self._free.put(k)
        self._copy_stream = torch.cuda.Stream(device=self.device) if self.cuda else None
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._thread.start()

    # ---- producer thread
    def _batches(self) -> Iterator[np.ndarray]:
        first = True
        while True:
            order = epoch_indices(len(self.ds), self.sampler, batch_size=self.B, seed=self.seed,
                                  start_idx=self.start_idx if (first or self.reapply_start_idx) else 0, rng=self._rng)
            first = False
            for k in range(len(order) // self.B):  # drop_last
                yield order[k * self.B: (k + 1) * self.B]

    def _run(self) -> None:
        try:
            if self.cuda:
                torch.cuda.set_device(self.device)
            for idx in self._batches():
                slot = None
                while slot is None:
                    if self._stop.is_set():
                        return
                    try:
                        slot = self._free.get(timeout=0.05)
                    except queue.Empty:
                        continue
                host, dev = self._slots[slot]
                batch = self.ds.rows(idx)
                inputs, labels, doc_start = prepare_batch(batch, self.T, self.use_doc_mask, self.vocab_size)
                # The slot's previous host -> device copy may still be QUEUED (the copy stream waits for the compute
                # stream at every release): the pinned buffers must not be rewritten before it has actually run.
                # Waiting here costs nothing: this is the worker thread.
                if self._slot_copy_done[slot] is not None:
                    self._slot_copy_done[slot].synchronize()
                host[0].copy_(inputs)
                host[1].copy_(labels)
                if self.use_doc_mask:
                    host[2].copy_(doc_start)
                ready = None
                if self.cuda:
                    with torch.cuda.stream(self._copy_stream):
                        for h, d in zip(host, dev):
                            d.copy_(h, non_blocking=True)
                        ready = torch.cuda.Event()
                        ready.record(self._copy_stream)
                    self._slot_copy_done[slot] = ready
                item = DeviceBatch(dev[0], dev[1], dev[2] if self.use_doc_mask else None, ready, slot)
                while not self._stop.is_set():
                    try:
                        self._q.put(item, timeout=0.05)
                        break
                    except queue.Full:
                        continue
        except BaseException as e:  # surfaced to the consumer
            self._err = e
            self._q.put(None)

    # ---- consumer side
    def __iter__(self) -> "BatchFeeder":
        return self

    def __next__(self) -> DeviceBatch:
        item = self._q.get()
        if item is None:
            raise RuntimeError("BatchFeeder producer failed") from self._err
        if item.ready is not None:
            torch.cuda.current_stream(self.device).wait_event(item.ready)
        self.consumed += 1
        return item

    def release(self, item: DeviceBatch) -> None:
        """Hand a batch's buffers back once the step that reads them has been *queued* (stream order protects the
        device tensors: the refill's copy is issued on the copy stream after an event recorded here)."""
        if self.cuda:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.device))
            self._copy_stream.wait_event(ev)
        self._free.put(item._slot)

    def state(self) -> Dict[str, int]:
        per_epoch = len(self.ds) // self.B
        total = self.consumed
        if self.reapply_start_idx and self.start_idx:
            # every epoch of this stream is the tail [start_idx, per_epoch) (the reference's behaviour after a resume)
            run, tail = total - self.start_idx, per_epoch - self.start_idx
            return {"epoch": run // tail, "start_idx": self.start_idx + run % tail, "micro_batches": total}
        return {"epoch": total // per_epoch, "start_idx": total % per_epoch, "micro_batches": total}

    def close(self) -> None:
        self._stop.set()
        try:
            while True:
                self._q.get_nowait()
        except queue.Empty:
            pass
        self._thread.join(timeout=2.0)

    def __del__(self):
        try:
            self._stop.set()
        except Exception:
            pass
