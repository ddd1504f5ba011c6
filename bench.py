#!/usr/bin/env python
"""bench.py — train tokens/s (fwd + bwd [+ data-parallel gradient all-reduce]) of the models/LM causal Transformer.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config c3] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

Workload (BASELINE.json metric, config C3): the 420M LM of config/tr_420M_x8gpu.yaml (24 layers, d=1024, 16 heads,
hid=2730, vocab 50280, seq 2048), per-GPU micro-batch 8 x 2048 tokens, bf16 compute / fp32 masters, synthetic
uniform tokens, random-init weights.  One "step" = zero grads, bf16 weight refresh, forward, backward and (N>1) the
bucketed gradient all-reduce.  Prints ONE JSON line on rank 0.

`value`  : tokens/s with the token ids already resident in HBM (CUDA events, max over ranks).
`e2e`    : the same through the public API with HOST rows: BatchFeeder (collate + _prepare_batch on a worker thread,
           pinned ring, H2D on a copy stream) -> Engine.compute_grads, H2D copies and a D2H read of the loss inside the
           timed region every step.
`roofline`: the dominant kernel (the tcgen05 GEMM): model GEMM FLOPs of one step / the CUDA-event time of exactly that
           GEMM sequence replayed back to back, against MEASURED_PEAKS.json's sustained bf16 figure.
`cpu_baseline` / `--impl reference`: the reference's JAX path cannot run here (jax is not installed); the timed CPU
           arm is the oracle port (oracle/lm_oracle.py, plain PyTorch fp32) on the host cores, on a bounded sample.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

CONFIGS = {
    # name: (n_layers, d_model, n_heads, vocab, seq_len, per-GPU micro-batch, doc-masked)
    "c1": dict(n_layers=4, d_model=256, n_heads=4, vocab_size=50257, seq_len=256, micro_batch=8, doc=False),
    "c2": dict(n_layers=12, d_model=768, n_heads=12, vocab_size=50257, seq_len=1024, micro_batch=16, doc=False),
    "c3": dict(n_layers=24, d_model=1024, n_heads=16, vocab_size=50280, seq_len=2048, micro_batch=8, doc=False),
    "c4": dict(n_layers=24, d_model=1024, n_heads=16, vocab_size=50280, seq_len=4096, micro_batch=4, doc=False),
    "c5": dict(n_layers=12, d_model=768, n_heads=12, vocab_size=50280, seq_len=2048, micro_batch=8, doc=True),
}
WORKLOAD_NAMES = {
    "c1": "tiny LM 4L d256 seq256",
    "c2": "GPT-2-small-sized LM 12L d768 seq1024",
    "c3": "420M LM config/tr_420M_x8gpu.yaml 24L d1024 seq2048",
    "c4": "420M LM long-context seq4096",
    "c5": "doc-masked LM config/config_doc_mask.yaml 12L d768 seq2048",
}


def yaml_like_cfg(c: dict) -> dict:
    """The reference's YAML keys (config/tr_420M_x8gpu.yaml) for construct_model."""
    return dict(model="transformer", vocab_size=c["vocab_size"], d_model=c["d_model"], expand="8/3",
                n_layers=c["n_layers"], n_heads=c["n_heads"], mlp_class="glu", seq_len=c["seq_len"],
                tie_embeddings=False, dtype="bfloat16", seed=100)


def flops_per_token(c: dict) -> float:
    """6*(L*(4d^2 + 3*d*hid) + d*V) + 6*L*T*d  (un-padded dims, causal attention at half)."""
    d, L, V, T = c["d_model"], c["n_layers"], c["vocab_size"], c["seq_len"]
    hid = int(8 / 3 * d)
    return 6.0 * (L * (4 * d * d + 3 * d * hid) + d * V) + 6.0 * L * T * d


def gemm_flops_per_token(c: dict) -> float:
    d, L, V = c["d_model"], c["n_layers"], c["vocab_size"]
    hid = int(8 / 3 * d)
    return 6.0 * (L * (4 * d * d + 3 * d * hid) + d * V)


def load_gemm_traffic(cname: str):
    """DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) of one step's GEMM launches, from the committed
    `ncu --set full` capture summarised by tools/gemm_traffic.py (C3 shapes only); None when there is no capture."""
    if cname != "c3":
        return None
    for name in ("r02_gemm_traffic.json", "r01_gemm_traffic.json"):
        p = os.path.join(ROOT, "profiles", name)
        if os.path.exists(p):
            with open(p) as f:
                return json.load(f).get("step_bytes")
    return None


def load_peaks() -> dict:
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        d["_source"] = "measured (MEASURED_PEAKS.json)"
        return d
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0,
            "_source": "fallback (B200_PROFILING.md)"}


class ClockSampler:
    """Samples nvidia-smi SM clocks / throttle reasons for one GPU while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.rows = []
        self._stop = threading.Event()
        self._t = None

    def _run_nvml(self) -> bool:
        """Preferred: NVML in-process (a sample costs well under a millisecond; eight ranks forking nvidia-smi at once
        got two samples into a one-second timed region).  False if NVML cannot be used: the caller forks nvidia-smi."""
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            smax = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
                nv.nvmlDeviceGetCurrentClocksThrottleReasons
            bits = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20),
                    ("sw_power_cap", 0x4))
            float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))  # (raises here rather than inside the loop)
        except Exception:
            return False
        while not self._stop.is_set():
            try:
                sm = float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                mask = int(get_reasons(h))
                watts = nv.nvmlDeviceGetPowerUsage(h) / 1000.0
                self.rows.append([str(sm), str(smax), str(watts)] + [("Active" if mask & b else "Not Active") for _, b in bits])
            except Exception:
                pass
            self._stop.wait(0.05)
        return True

    def _run(self):
        if self._run_nvml():
            return
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.2)

    def start(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()

    def stop(self) -> dict:
        self._stop.set()
        if self._t is not None:
            self._t.join(timeout=6)
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                smax.append(float(r[1]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------ CPU arm (oracle)
def run_cpu_arm(cname: str, steps: int, warmup: int, max_seconds: float = 150.0) -> dict:
    """Times the oracle port (the only runnable form of the reference's CPU path here) on the host cores:
    fwd + bwd of the same model on a bounded sample (ONE sequence of the config), all host threads."""
    import torch
    from oracle import lm_oracle as O
    c = CONFIGS[cname]
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg = O.ModelConfig(vocab_size=c["vocab_size"], seq_len=c["seq_len"], dim=c["d_model"], expand=8 / 3,
                        n_layers=c["n_layers"], n_heads=c["n_heads"])
    params = O.init_params(cfg, seed=0)
    B = 1
    batch = O.synthetic_batch(B, c["seq_len"], c["vocab_size"], seed=1234, doc_mean=512 if c["doc"] else None)
    inputs, labels, mask = O.prepare_batch(batch, c["seq_len"], c["doc"])
    times = []
    t_begin = time.perf_counter()
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        O.compute_grads(params, cfg, inputs, labels, mask, torch.float32)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
        if time.perf_counter() - t_begin > max_seconds and len(times) >= 1:
            break
    ms = 1e3 * sum(times) / len(times)
    tokens = B * c["seq_len"]
    return {"value": tokens / (ms / 1e3), "unit": "tokens/s", "cores": cores, "kind": "port",
            "sample": f"oracle/lm_oracle.py fp32 fwd+bwd, {B} sequence x {c['seq_len']} tokens of {cname}, "
                      f"{len(times)} timed step(s) after {warmup} warm-up",
            "ms_per_step": ms, "steps": len(times)}


def reference_main(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    c = CONFIGS[args.config]
    cb = run_cpu_arm(args.config, max(1, min(args.steps, 5)), max(0, min(args.warmup, 1)))
    line = {
        "impl": "reference", "metric": "train tokens/s fwd+bwd", "value": cb["value"], "unit": "tokens/s",
        "n_gpus": args.gpus, "steps": cb["steps"], "warmup": max(0, min(args.warmup, 1)),
        "ms_per_step": cb["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD_NAMES[args.config], "sample": cb["sample"],
                   "note": "reference is JAX/Flax; jax is not installable here, so this arm times the oracle port "
                           "of its CPU path on the host cores"},
        "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": cb["value"], "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ GPU arm
def gemm_replay_ms(model, M: int, iters: int = 3) -> float:
    """CUDA-event time of exactly the GEMM launches of one training step (fwd, dgrad, wgrad of every Dense),
    replayed back to back on scratch buffers; returns ms per step-worth of GEMMs."""
    import torch
    from plaincv_b200 import ops
    L = model.layout
    d, hp, Vp = L.d, L.hid_pad, L.V_pad
    dev = model.flat_grad.device
    bf = torch.bfloat16
    x_d = torch.randn(M, d, device=dev).to(bf)
    x_hp = torch.randn(M, hp, device=dev).to(bf)
    y_3d = torch.randn(M, 3 * d, device=dev).to(bf)
    y_2hp = torch.randn(M, 2 * hp, device=dev).to(bf)
    y_V = torch.randn(M, Vp, device=dev).to(bf) * 0.01
    o_d = torch.empty(M, d, device=dev, dtype=bf)
    o_hp = torch.empty(M, hp, device=dev, dtype=bf)
    layer = model.layers[0]
    gscratch = torch.zeros_like(model.flat_grad)
    g = lambda name: L.slot_view(gscratch, name)  # noqa: E731

    def lin(x, w, y):
        Mx, K = x.shape
        Np = w.shape[1]
        ops.gemm(x, w, y, Mx, Np, K, K, Np, Np, False, True, ops.PCV_EPI_BF16)

    def dgrad(dy, w, dx):
        Mx, Np = dy.shape
        K = w.shape[0]
        ops.gemm(dy, w, dx, Mx, K, Np, Np, Np, K, False, False, ops.PCV_EPI_BF16)

    def one_step():
        for _ in range(model.cfg.n_layers):
            lin(x_d, layer.w_qkv, y_3d)
            lin(x_d, layer.w_out, o_d)
            ops.linear_fwd_swiglu(x_d, layer.w_gu)            # gate|up GEMM, SwiGLU in its epilogue
            lin(x_hp, layer.w_down, o_d)
            ops.linear_dgrad_dswiglu(x_d, layer.w_down, y_2hp)  # fc2 dgrad, SwiGLU backward in its epilogue
            ops.linear_wgrad(x_hp, x_d, g("layers_0.w_down"))
            dgrad(y_2hp, layer.w_gu, o_d)
            ops.linear_wgrad(x_d, y_2hp, g("layers_0.w_gu"))
            dgrad(x_d, layer.w_out, o_d)
            ops.linear_wgrad(x_d, x_d, g("layers_0.w_out"))
            dgrad(y_3d, layer.w_qkv, o_d)
            ops.linear_wgrad(x_d, y_3d, g("layers_0.w_qkv"))
        ops.head_logits_ce(x_d, model.head)               # LM head with the softmax partials in its epilogue
        dgrad(y_V, model.head.w_lm, o_d)
        ops.linear_wgrad(x_d, y_V, g("w_lm"))

    one_step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        one_step()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def attn_flops_per_token(c: dict, docs_lengths=None) -> float:
    """Attention term of BASELINE.md §3.  Plain causal: 6*L*T*d per token.  Document-masked: every document of length n
    attends n^2/2 pairs instead of T^2/2, so the term becomes 6*L*d*sum(n_i^2)/T, averaged over the batch's rows (the
    lengths of the inputs, i.e. after train_lm.py:_trim_last_token)."""
    d, L, T = c["d_model"], c["n_layers"], c["seq_len"]
    if docs_lengths is None:
        return 6.0 * L * T * d
    from plaincv_b200.batch import trim_last_token
    sq = [sum(n * n for n in trim_last_token(row)) for row in docs_lengths]
    return 6.0 * L * d * (sum(sq) / len(sq)) / T


_KERNEL_CLASSES = (("gemm_bf16", "gemm"), ("attn_fwd", "attn_fwd"), ("attn_bwd", "attn_bwd"), ("rmsnorm", "rmsnorm"),
                   ("softmax_ce", "loss"), ("ce_merge", "loss"), ("ce_apply", "loss"), ("ce_finish", "loss"), ("mean2", "loss"), ("nccl", "nccl"), ("cast", "cast"),
                   ("embed", "embed"), ("swiglu", "swiglu"), ("Memset", "memset"), ("Memcpy", "memcpy"))


def profile_one_step(fn) -> dict:
    """Per-kernel durations INSIDE one real step (CUPTI through torch.profiler; the step's own streams, L2 state and
    side-stream overlap — unlike the replay).  Durations of kernels that overlap on different streams are summed per
    class, `busy_ms` is the union of all kernel intervals and `idle_ms` what is left of the step's span."""
    import torch
    from torch.profiler import ProfilerActivity, profile
    try:
        fn()
        torch.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
            fn()
            torch.cuda.synchronize()
        iv, by = [], {}
        for e in prof.events():
            if e.device_type != torch.autograd.DeviceType.CUDA:
                continue
            t0, t1 = e.time_range.start, e.time_range.end
            iv.append((t0, t1))
            cls = "other"
            for key, name in _KERNEL_CLASSES:
                if key in e.name:
                    cls = name
                    break
            rec = by.setdefault(cls, {"n": 0, "ms": 0.0})
            rec["n"] += 1
            rec["ms"] += (t1 - t0) / 1e3
        if not iv:
            return {"error": "profiler returned no device events"}
        iv.sort()
        busy, cur0, cur1 = 0.0, iv[0][0], iv[0][1]
        for a, b in iv[1:]:
            if a > cur1:
                busy += cur1 - cur0
                cur0, cur1 = a, b
            else:
                cur1 = max(cur1, b)
        busy += cur1 - cur0
        span = max(b for _, b in iv) - iv[0][0]
        return {"span_ms": span / 1e3, "busy_ms": busy / 1e3, "idle_ms": (span - busy) / 1e3,
                "by_kernel": {k: {"n": v["n"], "ms": round(v["ms"], 4)} for k, v in sorted(by.items())}}
    except Exception as ex:  # the profiler is evidence, not the product: never fail the bench on it
        return {"error": f"{type(ex).__name__}: {ex}"}


class Workload:
    """One BASELINE config on this rank's GPU: model, engine, resident inputs and the timed step functions."""

    def __init__(self, cname: str, dev, rank: int, world: int):
        import torch
        from plaincv_b200.batch import synthetic_batch
        from plaincv_b200.engine import Engine
        from plaincv_b200.model import construct_model
        self.cname, self.dev, self.rank, self.world = cname, dev, rank, world
        c = self.c = CONFIGS[cname]
        self.B, self.T, self.V = c["micro_batch"], c["seq_len"], c["vocab_size"]
        self.model, _ = construct_model(yaml_like_cfg(c), device=dev)
        # PCV_DDP_OVERLAP=0 (A/B runs only): all-reduce the buckets after backward instead of under it
        # PCV_DDP_COMM=fp32 (A/B runs only): all-reduce the gradient buckets in fp32 instead of bf16
        self.comm_fp32 = os.environ.get("PCV_DDP_COMM", "bf16") == "fp32"
        # PCV_CUDA_GRAPH=0 (A/B runs only): launch every kernel from Python instead of replaying the captured step
        self.cuda_graph = os.environ.get("PCV_CUDA_GRAPH", "1") != "0"
        self.eng = Engine(self.model, seq_len=self.T, use_doc_mask=c["doc"], cuda_graph=self.cuda_graph,
                          overlap=os.environ.get("PCV_DDP_OVERLAP", "1") != "0",
                          grad_comm_dtype=torch.float32 if self.comm_fp32 else torch.bfloat16)
        self.batch = synthetic_batch(self.B, self.T, self.V, seed=1234 + rank, doc_mean=512 if c["doc"] else None)
        inputs_h, labels_h, ds_h = eng_prepare(self.eng, self.batch)
        self.inputs_d, self.labels_d = inputs_h.to(dev), labels_h.to(dev)
        self.ds_d = None if ds_h is None else ds_h.to(dev)
        self.dev_batch = SimpleBatch(self.inputs_d, self.labels_d, self.ds_d)
        self.tokens_per_step = self.B * self.T * world
        self.fpt_causal = flops_per_token(c)
        self.fpt = gemm_flops_per_token(c) + attn_flops_per_token(c, self.batch.get("docs_lengths"))

    def step_resident(self):
        """bf16 weight refresh + Engine.compute_grads on the device-resident batch (zero grads, fwd, bwd, N > 1: the
        bucketed all-reduce under backward and the rank-mean of loss / acc)."""
        self.model.sync_compute_weights()
        loss, _ = self.eng.compute_grads(self.dev_batch)
        return loss

    def close(self):
        import torch
        self.model = self.eng = self.inputs_d = self.labels_d = self.ds_d = self.dev_batch = None
        torch.cuda.empty_cache()


def make_timer(dev, local_rank: int, world: int, lib):
    import torch
    import torch.distributed as dist

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def timed(fn, steps, warmup, eng=None):
        """launches = this library's kernels executed inside the timed region: direct C-ABI launches
        (pcv_launch_count) plus, when the engine replays its captured micro-step, the kernel nodes of every replay."""
        def count():
            return lib.pcv_launch_count() + (eng.replayed_launches if eng is not None else 0)

        for _ in range(warmup):
            fn()
        barrier()
        sampler = ClockSampler(local_rank)
        sampler.start()
        l0 = count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        launches = count() - l0
        clocks = sampler.stop()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms) / steps, launches, clocks

    return timed


def extra_config_line(cname: str, dev, rank, world, timed, peaks, min_seconds: float = 1.5) -> dict:
    """The other BASELINE configs (C2, C4, C5) at the same N, each timed after the headline: tokens/s, ms/step, MFU,
    clocks.  Enough steps for >= 3 clock samples (the 16 ms configs need ~100)."""
    w = Workload(cname, dev, rank, world)
    ms_probe, _, _ = timed(w.step_resident, 3, 3)
    steps = max(10, int(min_seconds * 1e3 / ms_probe) + 1)
    ms, launches, clocks = timed(w.step_resident, steps, 1, eng=w.eng)
    tok_s = w.tokens_per_step / (ms / 1e3)
    rec = {"workload": WORKLOAD_NAMES[cname], "value": tok_s, "unit": "tokens/s", "ms_per_step": ms, "steps": steps,
           "warmup": 7, "micro_batch_per_gpu": w.B, "seq_len": w.T, "n_gpus": world,
           "flop_per_token": w.fpt, "gpu_launches": int(launches), "clocks": clocks,
           "mfu": {"of_burst_peak": tok_s * w.fpt / (world * peaks["bf16_tflops"] * 1e12),
                   "of_sustained_peak": tok_s * w.fpt / (world * peaks["bf16_tflops_sustained"] * 1e12)}}
    if w.c["doc"]:
        rec["flop_per_token_plain_causal"] = w.fpt_causal
        rec["mfu"]["of_burst_peak_if_counted_as_plain_causal"] = tok_s * w.fpt_causal / (world * peaks["bf16_tflops"] * 1e12)
        rec["flops_note"] = "attention flops counted from this rank's actual documents: 6*L*d*sum(len^2)/T per token"
    w.close()
    return rec


def gpu_main(args) -> None:
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.gpus != world:
        raise SystemExit(f"bench.py: --gpus {args.gpus} but WORLD_SIZE={world}: launch N > 1 with "
                         "`python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N`")
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device; plaincv_b200 has no CPU fallback "
                           "(use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    from plaincv_b200 import _lib
    from plaincv_b200.batch import synthetic_batch

    lib = _lib.load()
    timed = make_timer(dev, local_rank, world, lib)
    w = Workload(args.config, dev, rank, world)
    c, B, T, V = w.c, w.B, w.T, w.V
    model, eng = w.model, w.eng
    comm_fp32 = w.comm_fp32
    tokens_per_step = w.tokens_per_step

    ms_step, launches, clocks = timed(w.step_resident, args.steps, args.warmup, eng=w.eng)

    # end to end through the public API: host rows -> BatchFeeder (collate + _prepare_batch on a worker thread, pinned
    # ring, H2D on a copy stream) -> Engine.compute_grads -> loss read back, every step
    from plaincv_b200.feed import BatchFeeder, InMemoryTokenDataset
    host_rows = InMemoryTokenDataset.synthetic(4 * B, T, V, seed=1234 + rank, doc_mean=512 if c["doc"] else None)
    feeder = BatchFeeder(host_rows, B, T, use_doc_mask=c["doc"], device=dev, depth=3, vocab_size=V)

    def step_e2e():
        model.sync_compute_weights()
        item = next(feeder)
        loss, acc = eng.compute_grads(item)
        feeder.release(item)
        return float(loss)  # device -> host read of the step's result

    ms_e2e, _, _ = timed(step_e2e, args.steps, min(args.warmup, 2) if args.warmup else 0)
    loss_val = step_e2e()
    feeder.close()

    # the reference's real schedule (config/tr_420M_x8gpu.yaml:35 grad_accumulation_steps: 4, train_lm.py:644-666): one
    # optimizer step = 4 micro-batches; here the all-reduce runs once, under the last micro-batch's backward
    micro = [{"input_ids": w.batch["input_ids"], **({"docs_lengths": w.batch["docs_lengths"]} if c["doc"] else {})}] * 4
    micro_dev = [SimpleBatch(w.inputs_d, w.labels_d, w.ds_d)] * 4

    def step_accum4():
        model.sync_compute_weights()
        return eng.train_step(micro_dev)

    del micro
    # 3 warm-up rounds: with CUDA graphs each (accumulate, sync) variant of the micro-step runs eagerly once and is
    # captured on its second call
    ms_acc4, _, _ = timed(step_accum4, max(2, args.steps // 4), 3)

    # (skipped under ncu: two CUPTI clients cannot share the device)
    in_step = {"skipped": True} if args.no_in_step_profile else profile_one_step(w.step_resident)

    ddp = None
    if world > 1 and not args.no_ddp_check:
        from plaincv_b200.selfcheck import ddp_gradient_check
        batches = [synthetic_batch(B, T, V, seed=1234 + k, doc_mean=512 if c["doc"] else None) for k in range(world)]
        ddp = ddp_gradient_check(model, T, c["doc"], batches, rank, world)

    line = None
    peaks = load_peaks()
    if rank == 0:
        fpt = w.fpt
        tok_s = tokens_per_step / (ms_step / 1e3)
        tok_s_e2e = tokens_per_step / (ms_e2e / 1e3)
        gemm_ms = gemm_replay_ms(model, B * T)
        gemm_flop = gemm_flops_per_token(c) * B * T
        gemm_tf = gemm_flop / (gemm_ms / 1e3) / 1e12
        if "by_kernel" in in_step and "gemm" in in_step["by_kernel"]:
            g_ms = in_step["by_kernel"]["gemm"]["ms"]
            in_step["gemm_tflops"] = gemm_flop / (g_ms / 1e3) / 1e12
            in_step["gemm_frac_of_sustained_peak"] = in_step["gemm_tflops"] / peaks["bf16_tflops_sustained"]
            in_step["note"] = ("kernel durations inside one real step (CUPTI); wgrad GEMMs run on a side stream and "
                               "overlap dgrad GEMMs, so class sums can exceed busy_ms")
        line = {
            "metric": "train tokens/s fwd+bwd", "value": tok_s, "unit": "tokens/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": WORKLOAD_NAMES[args.config], "global_batch": B * world, "seq_len": T,
                       "micro_batch_per_gpu": B, "parallelism": f"dp{world}", "params": model.count_params()[0],
                       "flop_per_token": fpt, "weights": "random-init (seed 100), fp32 masters + bf16 compute copies",
                       "l2": "working set per step (0.8 GB bf16 weights + ~13 GB activations at c3) >> 126 MB L2",
                       "step": "bf16 weight refresh + Engine.compute_grads (zero grads + fwd + bwd + (N>1) bucketed "
                               + ("fp32" if comm_fp32 else "bf16") + " gradient all-reduce under backward)"
                               + (", the micro-step replayed as one CUDA graph" if w.cuda_graph else "")},
            "mfu": {"of_burst_peak": tok_s * fpt / (world * peaks["bf16_tflops"] * 1e12),
                    "of_sustained_peak": tok_s * fpt / (world * peaks["bf16_tflops_sustained"] * 1e12),
                    "peaks": peaks["_source"]},
            "e2e": {"value": tok_s_e2e, "unit": "tokens/s", "h2d_bytes_per_step": int(2 * B * T * 4 +
                                                                                       (B * T * 4 if c["doc"] else 0)),
                    "d2h_bytes_per_step": 8, "ms_per_step": ms_e2e},
            "gpu_launches": int(launches),
            "gpu_launches_note": "this library's kernels executed in the timed region: direct C-ABI launches + the "
                                 "kernel nodes of every CUDA-graph replay of the micro-step",
            "clocks": clocks, "loss": loss_val,
            "roofline": {"kernel": "pcv::gemm_bf16_pair_kernel (all Dense fwd/dgrad/wgrad launches of one step)",
                         "bound": "tensor", "achieved": gemm_tf, "peak": peaks["bf16_tflops_sustained"] / 1.0,
                         "unit": "TFLOP/s", "frac": gemm_tf / peaks["bf16_tflops_sustained"],
                         "frac_of_burst_peak": gemm_tf / peaks["bf16_tflops"],
                         "traffic": load_gemm_traffic(args.config),
                         "traffic_note": "DRAM bytes of the same launches (one step), ncu --set full; "
                                         f"algorithmic flops per step = {gemm_flop:.4g}",
                         "peak_source": peaks["_source"] + " bf16_tflops_sustained",
                         "gemm_ms_per_step": gemm_ms, "gemm_share_of_step": gemm_ms / ms_step,
                         "in_step": in_step},
            "accum4": {"value": 4 * tokens_per_step / (ms_acc4 / 1e3), "unit": "tokens/s",
                       "ms_per_optimizer_step": ms_acc4,
                       "note": "the reference's schedule (grad_accumulation_steps: 4): 4 micro-batches per optimizer "
                               "step, one gradient all-reduce under the last backward"},
        }
        if ddp is not None:
            line["ddp_check"] = ddp
    w.close()
    del model, eng

    extras = {}
    if not args.no_extra_configs and args.config == "c3":
        for cname in ("c2", "c4", "c5"):
            rec = extra_config_line(cname, dev, rank, world, timed, peaks)
            if rank == 0:
                extras[cname] = rec
    if rank == 0:
        if extras:
            line["configs"] = extras
        if world == 1 and not args.no_cpu_baseline:
            cb = run_cpu_arm(args.config, 2, 1, max_seconds=90.0)
            line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
        print(json.dumps(line), flush=True)
    if world > 1:
        # No destroy_process_group(): tearing the NCCL communicator down while CUDA graphs that captured its collectives
        # still exist blocks forever (seen on 2 x B200, torch 2.11 / NCCL 2.28).  The line is printed and flushed; leave
        # without running destructors.
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


class SimpleBatch:
    """A prepared, device-resident micro-batch in the shape Engine accepts from feed.BatchFeeder."""

    def __init__(self, inputs, labels, doc_start):
        self.inputs, self.labels, self.doc_start = inputs, labels, doc_start


def eng_prepare(eng, batch):
    from plaincv_b200.batch import prepare_batch
    inputs, labels, ds = prepare_batch(batch, eng.seq_len, eng.use_doc_mask)
    return inputs.contiguous().pin_memory(), labels.contiguous().pin_memory(), (None if ds is None else ds.pin_memory())


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--config", default="c3", choices=sorted(CONFIGS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra-configs", action="store_true", help="skip the C2 / C4 / C5 lines appended under 'configs'")
    ap.add_argument("--no-in-step-profile", action="store_true", help="skip the CUPTI per-kernel profile of one step")
    ap.add_argument("--no-ddp-check", action="store_true", help="skip the N-rank gradient check run after the timed region")
    args = ap.parse_args()
    if args.impl == "reference":
        reference_main(args)
    else:
        gpu_main(args)


if __name__ == "__main__":
    main()
