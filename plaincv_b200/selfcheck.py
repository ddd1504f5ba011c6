"""On-hardware check of the data-parallel gradient path: N-rank gradients against the same global batch computed on
one rank (train_lm.py:151-159 shards the batch, :195-210 / :256-271 pmean the gradients — so the mean over ranks of the
shard gradients must equal the gradient of the whole batch).

Used by bench.py (after the timed region, whenever world > 1, so the driver's scaling runs carry the result) and by
tests/test_ddp_gpu.py (torchrun on 2+ GPUs).  Everything here runs the product path only: Engine, GradReducer, the
kernels; no oracle."""
from __future__ import annotations

from typing import Dict, Optional, Sequence

import torch
import torch.distributed as dist

from .engine import Engine


def _rel_l2(a: torch.Tensor, b: torch.Tensor) -> float:
    num = (a.double() - b.double()).norm()
    return float(num / (b.double().norm() + 1e-30))


@torch.no_grad()
def _snapshot(model) -> torch.Tensor:
    return model.flat_grad.clone()


def ddp_gradient_check(model, seq_len: int, use_doc_mask: bool, global_batches: Sequence[dict], rank: int, world: int,
                       process_group=None, layers_per_bucket: Optional[int] = None) -> Dict[str, object]:
    """`global_batches[k]` is rank k's micro-batch (every rank holds all of them: they are synthetic and seeded).

    1. the overlapped reducer with bf16 on the wire (the default), with fp32 on the wire, and with the overlap off
       (buckets reduced after backward) each produce rank-averaged gradients from this rank's shard;
    2. this rank alone accumulates all `world` micro-batches (sync=False, loss_scale 1/world): the gradient of the
       whole global batch — twice, to measure the run-to-run noise of the fp32 reduce-adds (split-K wgrad, dQ);
    3. relative L2 distances over the whole flat gradient buffer, worst rank reported (MAX all-reduce).
    Tolerances (asserted by the callers): bf16 wire <= 4e-3, fp32 wire <= max(1e-5, 2 x noise), overlap off vs on
    <= max(1e-5, 2 x noise) for the same wire format."""
    assert len(global_batches) == world
    mine = global_batches[rank]
    out: Dict[str, object] = {}

    def run(**kw):
        eng = Engine(model, seq_len=seq_len, use_doc_mask=use_doc_mask, process_group=process_group,
                     layers_per_bucket=layers_per_bucket, **kw)
        loss, _ = eng.compute_grads(mine)
        torch.cuda.synchronize()
        return _snapshot(model), float(loss)

    g_bf16, loss_ddp = run(grad_comm_dtype=torch.bfloat16, overlap=True)
    g_bf16_off, _ = run(grad_comm_dtype=torch.bfloat16, overlap=False)
    g_fp32, loss_ddp32 = run(grad_comm_dtype=torch.float32, overlap=True)
    g_fp32_off, _ = run(grad_comm_dtype=torch.float32, overlap=False)
    # the same step replayed as a CUDA graph (forward, backward, side-stream wgrads, casts and NCCL all-reduces of the
    # buckets on the communication stream, all captured): call 1 eager, call 2 captures, call 3 replays
    g_graph, loss_graph = None, float("nan")
    try:
        eng_g = Engine(model, seq_len=seq_len, use_doc_mask=use_doc_mask, process_group=process_group,
                       layers_per_bucket=layers_per_bucket, cuda_graph=True)
        for _ in range(3):
            lg, _ = eng_g.compute_grads(mine)
        torch.cuda.synchronize()
        g_graph, loss_graph = _snapshot(model), float(lg)
        graph_err = None
    except Exception as ex:  # reported, not fatal: the eager path is the reference for this check
        graph_err = f"{type(ex).__name__}: {ex}"

    # the whole global batch on this rank alone, no communication
    solo = Engine(model, seq_len=seq_len, use_doc_mask=use_doc_mask, process_group=process_group,
                  layers_per_bucket=layers_per_bucket)

    def whole_batch():
        stat = torch.zeros(2, dtype=torch.float32, device=model.flat_grad.device)
        for k, mb in enumerate(global_batches):
            solo._micro(mb, accumulate=k > 0, sync=False, loss_scale=1.0 / world, stat=stat)
        torch.cuda.synchronize()
        return _snapshot(model), float(stat[0]) / world

    g_ref, loss_ref = whole_batch()
    g_ref2, _ = whole_batch()

    vals = torch.tensor([
        _rel_l2(g_bf16, g_ref), _rel_l2(g_fp32, g_ref), _rel_l2(g_bf16_off, g_bf16), _rel_l2(g_fp32_off, g_fp32),
        _rel_l2(g_ref2, g_ref), abs(loss_ddp - loss_ref), abs(loss_ddp32 - loss_ref),
        _rel_l2(g_graph, g_bf16) if g_graph is not None else float("inf"), abs(loss_graph - loss_ddp),
    ], dtype=torch.float64, device=model.flat_grad.device)
    # ranks must also agree with each other bit for bit after the all-reduce: compare a checksum of the averaged buffer
    csum = g_bf16.double().sum().reshape(1)
    cmin, cmax = csum.clone(), csum.clone()
    if world > 1:
        dist.all_reduce(vals, op=dist.ReduceOp.MAX, group=process_group)
        dist.all_reduce(cmin, op=dist.ReduceOp.MIN, group=process_group)
        dist.all_reduce(cmax, op=dist.ReduceOp.MAX, group=process_group)
    v = [float(x) for x in vals]
    noise = v[4]
    out.update({
        "world": world, "micro_batches_per_rank": 1,
        "rel_l2_bf16_wire": v[0], "rel_l2_fp32_wire": v[1],
        "rel_l2_overlap_off_vs_on_bf16": v[2], "rel_l2_overlap_off_vs_on_fp32": v[3],
        "rel_l2_rerun_noise": noise,
        "loss_abs_diff": v[5], "loss_abs_diff_fp32_wire": v[6],
        "rel_l2_cuda_graph_vs_eager_bf16": v[7], "loss_abs_diff_cuda_graph": v[8], "cuda_graph_error": graph_err,
        "ranks_hold_identical_gradients": bool(float(cmin) == float(cmax)),
        "overlap_off_equal": bool(v[2] <= max(1e-5, 2.0 * noise) and v[3] <= max(1e-5, 2.0 * noise)),
        "ok": bool(v[0] <= 4e-3 and v[1] <= max(1e-5, 2.0 * noise) and v[5] <= 1e-4 and float(cmin) == float(cmax)),
        "tolerance": "bf16 wire <= 4e-3; fp32 wire and overlap on/off <= max(1e-5, 2 x rerun noise): two runs of the "
                     "same single-rank step differ by the order of fp32 reduce-adds (split-K wgrad, dQ), see "
                     "rel_l2_rerun_noise; loss |diff| <= 1e-4",
    })
    return out
