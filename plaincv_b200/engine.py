"""Step functions ("engine") for the LM hot path: the torch-side equivalent of train_lm.py's
compute_grads / eval_step (:189-313) and of its gradient-accumulation loop (:644-664), with the data-parallel
exchange of `jax.lax.pmean(grads, "data")` (:207, :268) done as a bucketed NCCL all-reduce that overlaps backward.

One process per GPU (torchrun).  Differences from the reference that are deliberate:
  * the all-reduce runs once per optimizer step (on the last micro-batch), not on every micro-step;
  * buckets are the per-layer slices of the model's flat fp32 gradient buffer, launched on a side stream as soon as a
    block's backward has finished (reverse layer order), so NVLink traffic hides under the remaining backward;
  * loss and accuracy are averaged in one extra 2-float all-reduce.
"""
from __future__ import annotations

import os

from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

from .batch import prepare_batch


class GradReducer:
    """Mean all-reduce of [begin, end) slices of one flat gradient buffer, one bucket at a time.

    Device-agnostic on purpose: with CUDA tensors the collectives are issued on a dedicated communication stream
    (fenced by events, so they overlap whatever the compute stream does next); with CPU tensors (gloo, used by the
    world_size=2 tests) they are issued as async work handles."""

    def __init__(self, flat_grad: torch.Tensor, ranges: Sequence[Tuple[int, int]], process_group=None,
                 comm_dtype: torch.dtype = torch.float32):
        """comm_dtype = torch.bfloat16 halves what crosses NVLink: a bucket is rounded to bf16 (round to nearest even),
        all-reduced in bf16 and written back to the fp32 buffer.  torch.float32 is the reference's pmean precision."""
        if comm_dtype not in (torch.float32, torch.bfloat16):
            raise ValueError("comm_dtype must be torch.float32 or torch.bfloat16")
        self.flat = flat_grad
        self.comm_dtype = comm_dtype
        self.ranges = list(ranges)
        self.pg = process_group
        self.world = dist.get_world_size(process_group) if dist.is_available() and dist.is_initialized() else 1
        self.cuda = flat_grad.is_cuda
        self.comm_stream = torch.cuda.Stream(device=flat_grad.device) if (self.cuda and self.world > 1) else None
        self._handles: List = []
        self._launched: List[int] = []
        backend = dist.get_backend(process_group) if self.world > 1 else None
        self._avg_op = dist.ReduceOp.AVG if backend == "nccl" else None
        self._wire = None
        if self.world > 1 and comm_dtype == torch.bfloat16 and self.cuda:
            self._wire = torch.empty(flat_grad.numel(), dtype=torch.bfloat16, device=flat_grad.device)

    def _all_reduce_mean(self, t: torch.Tensor, async_op: bool):
        if self._avg_op is not None:
            return dist.all_reduce(t, op=self._avg_op, group=self.pg, async_op=async_op)
        work = dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.pg, async_op=async_op)
        return (work, t)

    def bucket_ready(self, idx: int) -> None:
        """Call when every gradient inside bucket `idx` has been produced on the current stream."""
        if self.world == 1:
            return
        self._launched.append(idx)
        b, e = self.ranges[idx]
        view = self.flat[b:e]
        if self.cuda:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream())
            self.comm_stream.wait_event(ev)
            with torch.cuda.stream(self.comm_stream):
                if self._wire is not None:
                    from . import ops
                    wire = self._wire[b:e]
                    ops.cast_f32_bf16(view, wire)
                    self._all_reduce_mean(wire, async_op=False)
                    ops.cast_f32_bf16(wire, view)
                else:
                    self._all_reduce_mean(view, async_op=False)
        else:
            rounded = self.comm_dtype == torch.bfloat16
            if rounded:
                # CPU (gloo) runs of the host logic: same rounding points, the sum itself in fp32
                view.copy_(view.to(torch.bfloat16).to(torch.float32))
            self._handles.append((self._all_reduce_mean(view, async_op=True), view if rounded else None))

    def reduce_scalars(self, t: torch.Tensor) -> None:
        """Mean all-reduce of a small tensor (loss, acc) on the same path as the buckets."""
        if self.world == 1:
            return
        if self.cuda:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream())
            self.comm_stream.wait_event(ev)
            with torch.cuda.stream(self.comm_stream):
                self._all_reduce_mean(t, async_op=False)
        else:
            self._handles.append((self._all_reduce_mean(t, async_op=True), None))

    def finish_scalars(self) -> None:
        """Make the compute stream (or the host, for CPU tensors) wait for reduce_scalars."""
        if self.world == 1:
            return
        if self.cuda:
            torch.cuda.current_stream().wait_stream(self.comm_stream)
        else:
            self._wait_handles()

    def finish(self) -> None:
        """Launch any bucket not yet launched, then make the compute stream wait for all of them."""
        if self.world == 1:
            return
        for idx in range(len(self.ranges)):
            if idx not in self._launched:
                self.bucket_ready(idx)
        self._launched = []
        if self.cuda:
            torch.cuda.current_stream().wait_stream(self.comm_stream)
        else:
            self._wait_handles()

    def _wait_handles(self) -> None:
        for h, round_back in self._handles:
            if isinstance(h, tuple):
                work, t = h
                work.wait()
                t.div_(self.world)
            elif h is not None:
                h.wait()
            if round_back is not None:
                round_back.copy_(round_back.to(torch.bfloat16).to(torch.float32))
        self._handles = []


class _GraphedMicro:
    """One micro-step of an Engine captured as a CUDA graph (see Engine.__init__, cuda_graph).  Static int32 input
    buffers; every replay is preceded by three small device-to-device copies into them."""

    def __init__(self, eng: "Engine", key: tuple):
        self.eng = eng
        (self.shape, self.has_doc, self.accumulate, self.sync, self.loss_scale) = key
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        dev = eng.model.flat_grad.device
        self.inputs = torch.zeros(self.shape, dtype=torch.int32, device=dev)
        self.labels = torch.zeros(self.shape, dtype=torch.int32, device=dev)
        self.doc = torch.zeros(self.shape, dtype=torch.int32, device=dev) if self.has_doc else None
        self.stat = torch.zeros(2, dtype=torch.float32, device=dev)
        self.kernel_nodes = 0  # pcv_* launches recorded into the graph (one replay executes all of them)

    def _stage(self, inputs, labels, doc_start) -> None:
        self.inputs.copy_(inputs.view(self.shape), non_blocking=True)
        self.labels.copy_(labels.view(self.shape), non_blocking=True)
        if self.has_doc:
            self.doc.copy_(doc_start.view(self.shape), non_blocking=True)

    def run(self, inputs, labels, doc_start, stat: torch.Tensor) -> None:
        eng = self.eng
        self._stage(inputs, labels, doc_start)
        if self.graph is None:
            g = torch.cuda.CUDAGraph()
            if eng._graph_pool is None:
                eng._graph_pool = torch.cuda.graph_pool_handle()
            # thread_local: other host threads (the BatchFeeder's worker stages the next batches meanwhile) may keep
            # calling CUDA; autograd's device thread launches onto the capturing stream and is captured with it
            from . import _lib
            lib = _lib.load()
            n0 = lib.pcv_launch_count()
            with torch.cuda.graph(g, pool=eng._graph_pool, capture_error_mode="thread_local"):
                self.stat.zero_()
                eng._micro_body(self.inputs, self.labels, self.doc, self.accumulate, self.sync, self.loss_scale,
                                self.stat)
            self.kernel_nodes = int(lib.pcv_launch_count() - n0)
            self.graph = g
        self.graph.replay()
        eng.replayed_launches += self.kernel_nodes
        stat += self.stat


def bucket_groups(n_layers: int, layers_per_bucket: int) -> List[Tuple[int, int]]:
    """(lowest, highest) layer index of every gradient bucket, in the order backward completes them (top layers
    first).  Groups hold up to `layers_per_bucket` layers and taper — never more than half of the layers still ahead —
    down to single layers for the last two, so that little is left to reduce once backward has ended."""
    G = max(1, int(layers_per_bucket))
    groups = []
    hi = n_layers - 1
    while hi >= 0:
        size = min(G, max(1, (hi + 1) // 2))
        lo = hi - size + 1
        groups.append((lo, hi))
        hi = lo - 1
    return groups


class Engine:
    """compute_grads / eval_step / train_step over a plaincv_b200.Transformer."""

    def __init__(self, model, seq_len: Optional[int] = None, use_doc_mask: bool = False, process_group=None,
                 overlap: bool = True, grad_comm_dtype: torch.dtype = torch.bfloat16,
                 layers_per_bucket: Optional[int] = None, cuda_graph: Optional[bool] = None):
        """grad_comm_dtype: wire format of the gradient mean.  The default, bf16, is what BASELINE.json's north_star
        asks for ("bucketed bf16 gradient all-reduce") and deviates from the reference's fp32 `lax.pmean`
        (train_lm.py:207, 268) by the bf16 roundings of every bucket before, inside and after the reduction: the
        reducer alone, on known numbers, is within 2.45e-3 (2 GPUs) .. 4.04e-3 (8 GPUs) rel. L2 of the exact mean
        (bench.py `ddp_check`, tests/test_ddp_gpu.py) — below the run-to-run noise of the step's own fp32 reduce-adds
        at 24 layers (8e-3); torch.float32 reproduces the reference's precision.

        cuda_graph: replay one micro-step (zero grads, forward, backward, the side-stream wgrads and — N > 1 — the
        bucketed all-reduce on the communication stream) as ONE CUDA graph instead of ~490 kernel launches issued
        from Python: the first call of a given (batch shape, accumulate, sync, loss_scale) runs eagerly, the second
        is captured, later ones replay.  Same kernels, same order, same results; what goes away is the GPU idling
        between launches (about 2 ms of a 45 ms step at C3).  Default: env PCV_CUDA_GRAPH (off unless "1")."""
        self.model = model
        self.seq_len = int(seq_len if seq_len is not None else model.cfg.seq_len)
        self.use_doc_mask = bool(use_doc_mask)
        self.overlap = overlap
        self.pg = process_group
        self.world = dist.get_world_size(process_group) if dist.is_available() and dist.is_initialized() else 1
        L = model.layout
        # buckets in the order their gradients become final during backward: head, layers L-1..0 (consecutive layers
        # are contiguous in the flat buffer, so `layers_per_bucket` of them form one slice), embedding
        # Fewer, larger all-reduces cost less (26 -> 8 collectives at 24 layers: -0.85 ms per step on 2 GPUs), but what
        # the last bucket holds is reduced after backward has ended — so the groups taper: up to `layers_per_bucket`
        # (default 8) layers while much of backward is still ahead, halving towards single layers at the end.
        if layers_per_bucket is None:
            layers_per_bucket = int(os.environ.get("PCV_DDP_LAYERS_PER_BUCKET", "8"))
        self._groups = bucket_groups(len(L.layer_ranges), layers_per_bucket)
        self._ranges = ([L.head_range] + [(L.layer_ranges[lo][0], L.layer_ranges[hi][1]) for lo, hi in self._groups]
                        + [L.embed_range])
        self.reducer = GradReducer(model.flat_grad, self._ranges, process_group, comm_dtype=grad_comm_dtype)
        self._zero_stream = None
        self._zero_pending = False
        if cuda_graph is None:
            cuda_graph = os.environ.get("PCV_CUDA_GRAPH", "0") == "1"
        self.cuda_graph = bool(cuda_graph) and model.flat_grad.is_cuda
        self._graphs: Dict[tuple, "_GraphedMicro"] = {}
        self._graph_pool = None
        # pcv_* kernel launches executed through graph replays (pcv_launch_count() only sees direct C-ABI calls)
        self.replayed_launches = 0

    # ------------------------------------------------------------------ hooks
    def _arm_hooks(self, on: bool) -> None:
        m = self.model
        if on and self.world > 1 and self.overlap:
            m.head.on_grads_ready = lambda: self.reducer.bucket_ready(0)
            for layer in m.layers:
                layer.on_grads_ready = None
            for g, (lo, _hi) in enumerate(self._groups):  # a group is complete when its lowest layer's backward is
                m.layers[lo].on_grads_ready = (lambda k=1 + g: self.reducer.bucket_ready(k))
        else:
            m.head.on_grads_ready = None
            for layer in m.layers:
                layer.on_grads_ready = None

    # ------------------------------------------------------------------ steps
    def _to_device(self, batch):
        if hasattr(batch, "inputs") and hasattr(batch, "labels"):  # a feed.DeviceBatch: already prepared and staged
            if self.use_doc_mask and batch.doc_start is None:
                raise ValueError("Doc masking enabled but batch has no 'docs_lengths'.")  # train_lm.py:110
            return batch.inputs, batch.labels, (batch.doc_start if self.use_doc_mask else None)
        # host dicts are range-checked here (token ids index the embedding table and the logits unchecked in the kernels);
        # feed.BatchFeeder(vocab_size=...) does the same on its worker thread for staged batches
        inputs, labels, doc_start = prepare_batch(batch, self.seq_len, self.use_doc_mask, self.model.cfg.vocab_size)
        dev = self.model.flat_grad.device

        def up(t):
            if t is None:
                return None
            t = t.contiguous()
            if not t.is_pinned():
                t = t.pin_memory()
            return t.to(dev, non_blocking=True)

        return up(inputs), up(labels), up(doc_start)

    def zero_grads_async(self) -> None:
        """Zero the flat gradient buffer.  PCV_ZERO_ASYNC=1 moves the 1.6 GB memset (0.25 ms at C3) to a side stream
        so that it runs under the forward pass (nothing touches the buffer before backward) — measured on B200 this
        LOSES 1 % (47.9 vs 47.4 ms per step: the memset takes HBM / L2 bandwidth from the forward GEMMs for longer than
        it would have run alone), so the default zeroes in stream order."""
        g = self.model.flat_grad
        if not g.is_cuda or os.environ.get("PCV_ZERO_ASYNC", "0") != "1":
            g.zero_()
            return
        if self._zero_stream is None:
            self._zero_stream = torch.cuda.Stream(device=g.device)
        self._zero_stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self._zero_stream):
            g.zero_()
        self._zero_pending = True

    def wait_zero_grads(self) -> None:
        """Backward (the first writer of the gradient buffer) must not start before the zero-fill has finished."""
        if self._zero_pending:
            torch.cuda.current_stream().wait_stream(self._zero_stream)
            self._zero_pending = False

    def _micro_body(self, inputs, labels, doc_start, accumulate: bool, sync: bool, loss_scale: float,
                    stat: torch.Tensor) -> None:
        """One micro-step on device tensors; `stat` (2 floats) receives += this rank's (loss, acc); the gradient buckets
        are rank-averaged when sync (the two scalars are averaged by the caller, _reduce_stat, outside any graph)."""
        m = self.model
        if not accumulate:
            self.zero_grads_async()
        self._arm_hooks(sync)
        loss, acc = m.loss_and_acc(inputs, labels, doc_start=doc_start, loss_scale=loss_scale, unit_grad=True)
        self.wait_zero_grads()
        loss.backward()
        stat[0] += loss.detach()
        stat[1] += acc.detach()
        if sync and self.world > 1:
            self.reducer.finish()
        self._arm_hooks(False)

    def _reduce_stat(self, stat: torch.Tensor) -> None:
        """Rank-mean of the (loss, acc) sums (the pmean of train_lm.py:268-270 applied to the two scalars)."""
        if self.world > 1:
            self.reducer.reduce_scalars(stat)
            self.reducer.finish_scalars()

    def _micro(self, batch: dict, accumulate: bool, sync: bool, loss_scale: float, stat: torch.Tensor) -> None:
        inputs, labels, doc_start = self._to_device(batch)
        if not self.cuda_graph:
            self._micro_body(inputs, labels, doc_start, accumulate, sync, loss_scale, stat)
            return
        key = (tuple(inputs.shape), doc_start is not None, bool(accumulate), bool(sync), float(loss_scale))
        g = self._graphs.get(key)
        if g is None:  # first call of this kind: eager (it is also the warm-up CUDA graphs need before a capture)
            self._graphs[key] = _GraphedMicro(self, key)
            self._micro_body(inputs, labels, doc_start, accumulate, sync, loss_scale, stat)
            return
        g.run(inputs, labels, doc_start, stat)

    def compute_grads(self, batch: dict, *, accumulate: bool = False, sync: bool = True, loss_scale: float = 1.0):
        """One micro-batch of fwd + bwd (train_lm.py:256-271).  Gradients land in model.flat_grad (param.grad are views):
        zeroed first unless `accumulate`; averaged over ranks when `sync` (with the all-reduce overlapping backward).
        Returns (loss, acc) as 0-dim device tensors, rank-averaged when `sync`."""
        stat = torch.zeros(2, dtype=torch.float32, device=self.model.flat_grad.device)
        self._micro(batch, accumulate, sync, loss_scale, stat)
        if sync:
            self._reduce_stat(stat)
        return stat[0], stat[1]

    def train_step(self, micro_batches: Sequence[dict]):
        """The accumulation loop of train_lm.py:644-664 for one optimizer step: mean of the micro-batch gradients,
        rank-averaged — with a single all-reduce (issued during the last micro-batch's backward), where the reference
        all-reduces on every micro-step.  Returns the mean (loss, acc) over micro-batches and ranks."""
        n = len(micro_batches)
        stat = torch.zeros(2, dtype=torch.float32, device=self.model.flat_grad.device)
        for k, mb in enumerate(micro_batches):
            self._micro(mb, accumulate=k > 0, sync=(k == n - 1), loss_scale=1.0 / n, stat=stat)
        self._reduce_stat(stat)
        return stat[0] / n, stat[1] / n

    @torch.no_grad()
    def eval_step(self, batch: dict):
        """train_lm.py:273-284: forward + loss/acc, rank-averaged."""
        inputs, labels, doc_start = self._to_device(batch)
        loss, acc = self.model.loss_and_acc(inputs, labels, doc_start=doc_start)
        stat = torch.stack([loss, acc])
        if self.world > 1:
            dist.all_reduce(stat, op=dist.ReduceOp.SUM, group=self.pg)
            stat = stat / self.world
        return stat[0], stat[1]

    def apply_grads(self, optimizer, lr: Optional[float] = None) -> None:
        """train_lm.py:316-353: clip + optimizer update from the (already averaged) gradients of the last
        compute_grads / train_step; `optimizer` is a plaincv_b200.optim.AdamW.  The bf16 compute copies are
        refreshed by the same kernel."""
        optimizer.step(lr=lr)

    def after_optimizer_step(self) -> None:
        """Refresh the bf16 compute copies of the parameters (one cast kernel over the flat buffer)."""
        self.model.sync_compute_weights()

    def grads(self) -> Dict[str, torch.Tensor]:
        return self.model.export_flax_grads()

    def close(self) -> None:
        """Drop the captured CUDA graphs.  Call before torch.distributed.destroy_process_group(): a communicator whose
        collectives live on inside a captured graph does not tear down (the call blocks)."""
        self._graphs.clear()
        self._graph_pool = None
