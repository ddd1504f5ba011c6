#!/usr/bin/env python
"""Generates tests/golden/*.json from the CPU oracle (oracle/lm_oracle.py, fp32 mode).

The reference (JAX/Flax) cannot be imported in this environment and ships no golden vectors of its own, so these
fixtures pin the ORACLE's outputs (loss, accuracy, per-leaf gradient norms and 8 sampled gradient entries per leaf)
for seeded inputs; the CPU suite re-derives them from the oracle (guards against oracle drift) and the GPU suite holds
the CUDA path to them without running the oracle.  Parity with the reference itself stays "unpinned" (DESIGN.md §3).

usage: python tools/make_golden.py            (rewrites tests/golden/*.json; deterministic)
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from oracle import lm_oracle as O  # noqa: E402

CASES = {
    # BASELINE config C1 (tiny LM: 4 layers, d=256, 4 heads, seq 256, GPT-2 vocab), batch 8, plain causal
    "c1_causal": dict(vocab_size=50257, seq_len=256, dim=256, n_layers=4, n_heads=4, batch=8, doc_mean=None, seed=0,
                      data_seed=1234),
    # same model, packed documents (config_doc_mask.yaml's intra_doc_masking), batch 4
    "c1_docmask": dict(vocab_size=50257, seq_len=256, dim=256, n_layers=4, n_heads=4, batch=4, doc_mean=60, seed=1,
                       data_seed=77),
    # a 2-layer, 2-head model with a small ragged vocab: cheap enough for every CPU run
    "small_ragged": dict(vocab_size=1003, seq_len=128, dim=128, n_layers=2, n_heads=2, batch=3, doc_mean=40, seed=2,
                         data_seed=5, ggn_seed=17),
}


def sample_index(name: str, numel: int, k: int = 8):
    rng = np.random.default_rng(abs(hash_name(name)) % (2 ** 32))
    return sorted(int(i) for i in rng.choice(numel, size=min(k, numel), replace=False))


def hash_name(name: str) -> int:
    h = 2166136261
    for ch in name.encode():
        h = ((h ^ ch) * 16777619) & 0xFFFFFFFF
    return h


def ggn_direction(params, seed: int):
    """The seeded parameter-shaped direction of the GGN fixtures (kernels N(0, 0.02), norm scales N(0, 0.1))."""
    g = torch.Generator().manual_seed(seed)
    return {k: torch.randn(v.shape, generator=g) * (0.02 if v.dim() > 1 else 0.1) for k, v in params.items()}


def run_case(spec: dict) -> dict:
    cfg = O.ModelConfig(vocab_size=spec["vocab_size"], seq_len=spec["seq_len"], dim=spec["dim"], expand=8 / 3,
                        n_layers=spec["n_layers"], n_heads=spec["n_heads"])
    params = O.init_params(cfg, seed=spec["seed"])
    batch = O.synthetic_batch(spec["batch"], spec["seq_len"], spec["vocab_size"], seed=spec["data_seed"],
                              doc_mean=spec["doc_mean"])
    use_doc = spec["doc_mean"] is not None
    inputs, labels, mask = O.prepare_batch(batch, spec["seq_len"], use_doc)
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    grads, loss, acc = O.compute_grads(params, cfg, inputs, labels, mask, torch.float32)
    leaves = {}
    for k, g in grads.items():
        flat = g.reshape(-1)
        idx = sample_index(k, flat.numel())
        leaves[k] = {"shape": list(g.shape), "norm": float(g.double().norm()), "index": idx,
                     "values": [float(flat[i]) for i in idx]}
    extra = {}
    if spec.get("ggn_seed") is not None:
        # GGN matvec (optim/ggn_utils.py:69-153) along a seeded direction: per-leaf norms, sampled entries
        vec = ggn_direction(params, spec["ggn_seed"])
        gv = O.ggn_matvec(params, cfg, vec, inputs, labels, mask, torch.float32)
        rec = {}
        for k, g in gv.items():
            flat = g.reshape(-1)
            idx = sample_index(k, flat.numel())
            rec[k] = {"norm": float(g.double().norm()), "index": idx, "values": [float(flat[i]) for i in idx]}
        extra["ggn"] = {"seed": spec["ggn_seed"], "vGv": float(sum((vec[k].double() * gv[k].double()).sum() for k in gv)),
                        "leaves": rec}
    return {"spec": spec, "loss": float(loss), "acc": float(acc), "n_params": O.count_params(params), **extra,
            "input_ids_sum": int(np.asarray(batch["input_ids"]).astype(np.int64).sum()),
            "docs_lengths": batch.get("docs_lengths"), "grads": leaves}


def main():
    out_dir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    for name, spec in CASES.items():
        res = run_case(spec)
        with open(os.path.join(out_dir, f"{name}.json"), "w") as f:
            json.dump(res, f, indent=1)
        print(name, "loss", res["loss"], "acc", res["acc"], "params", res["n_params"])


if __name__ == "__main__":
    main()
