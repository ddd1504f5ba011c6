"""CPU suite, part 2: host logic of the package (batch preparation, parameter layout, error contract), the C-ABI
library (loads, exports every symbol include/plaincv_b200.h declares — no compute calls without a GPU) and the
data-parallel gradient path over gloo with world_size 2."""
import ctypes
import os
import re
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

from oracle import lm_oracle as O  # noqa: E402  (checker only)
from plaincv_b200 import batch as PB  # noqa: E402


# ------------------------------------------------------------------------------------------------ batch contract
def test_prepare_batch_matches_reference_contract():
    b = PB.synthetic_batch(3, 16, 101, seed=5, doc_mean=5)
    inputs, labels, ds = PB.prepare_batch(b, 16, True)
    ids = torch.as_tensor(b["input_ids"])
    assert torch.equal(inputs, ids[:, :-1]) and torch.equal(labels, ids[:, 1:])
    assert inputs.dtype == torch.int32 and tuple(ds.shape) == (3, 16)
    # doc_start carries exactly the mask train_lm.py builds (oracle restates _build_attn_mask)
    mask = O.build_attn_mask(b, 16)
    assert np.array_equal(PB.dense_mask_from_doc_start(ds.numpy()), mask.numpy())
    assert torch.equal(PB.doc_start_from_mask(mask), ds)
    # same generator as the oracle's (bench arms see identical tokens)
    bo = O.synthetic_batch(3, 16, 101, seed=5, doc_mean=5)
    assert np.array_equal(bo["input_ids"], b["input_ids"]) and bo["docs_lengths"] == b["docs_lengths"]


def test_prepare_batch_errors():
    with pytest.raises(ValueError, match="seq_len\\+1"):
        PB.prepare_batch({"input_ids": np.zeros((2, 16), np.int32)}, 16, False)
    with pytest.raises(ValueError, match="docs_lengths not found"):
        PB.prepare_batch({"input_ids": np.zeros((2, 17), np.int32)}, 16, True)
    with pytest.raises(ValueError, match="Sum\\(doc_boundaries\\)"):
        PB.prepare_batch({"input_ids": np.zeros((1, 17), np.int32), "docs_lengths": [[5, 5]]}, 16, True)
    with pytest.raises(ValueError):
        PB.prepare_batch({"input_ids": np.zeros((2, 17), np.int32), "docs_lengths": [[17]]}, 16, True)


def test_trim_last_token_edge_cases():
    assert PB.trim_last_token([4, 1]) == [4]          # a trailing 1-token document vanishes
    assert PB.trim_last_token([5]) == [4]
    assert PB.trim_last_token([]) == []
    ds = PB.doc_lengths_to_doc_start([[4, 1], [1, 1, 3]], 4)
    assert ds.tolist() == [[0, 0, 0, 0], [0, 1, 2, 2]]


def test_doc_start_from_mask_rejects_other_masks():
    T = 6
    full = torch.ones(1, T, T, dtype=torch.bool)
    with pytest.raises(ValueError):
        PB.doc_start_from_mask(full)                  # not causal
    with pytest.raises(ValueError):
        PB.doc_start_from_mask(torch.zeros(1, T, T))  # additive / float mask
    causal = torch.tril(torch.ones(T, T, dtype=torch.bool))[None]
    assert PB.doc_start_from_mask(causal).tolist() == [[0] * T]


def test_shard_for_rank_is_pmap_split():
    x = torch.arange(24).view(8, 3)
    assert torch.equal(PB.shard_for_rank(x, 1, 4), x[2:4])
    with pytest.raises(ValueError, match="not divisible"):
        PB.shard_for_rank(x, 0, 3)
    assert PB.shard_for_rank(None, 0, 2) is None


# ------------------------------------------------------------------------------------------------ parameter layout
def test_param_layout_mirrors_flax_tree():
    from plaincv_b200.model import ModelConfig, ParamLayout, model_config_from_cfg
    cfg = model_config_from_cfg(dict(model="transformer", vocab_size=50257, d_model=256, expand="8/3", n_layers=4,
                                     n_heads=4, mlp_class="glu", seq_len=256, dtype="bfloat16"))
    assert cfg.hidden_dim == 682
    L = ParamLayout(cfg)
    ocfg = O.ModelConfig(vocab_size=50257, seq_len=256, dim=256, expand=8 / 3, n_layers=4, n_heads=4)
    shapes = O.param_shapes(ocfg)
    assert {v.name: tuple(v.shape) for v in L.views} == shapes
    # every padded activation row stride is a whole number of 128-byte lines in bf16; slots start 256-byte aligned
    assert L.hid_pad % 64 == 0 and L.V_pad % 64 == 0 and L.hid_pad >= 682 and L.V_pad >= 50257
    for s in L.slots.values():
        assert (s.ld * 2) % 16 == 0 and (s.offset * 2) % 256 == 0
    # views of distinct parameters never overlap, and pad lanes belong to nobody
    flat = torch.zeros(L.total, dtype=torch.float32)
    for v in L.views:
        L.param_view(flat, v).add_(1.0)
    assert float(flat.max()) == 1.0
    assert int(flat.sum()) == sum(int(np.prod(s)) for s in shapes.values())
    # buckets tile the buffer in order: embedding, layers, head
    ranges = [L.embed_range] + L.layer_ranges + [L.head_range]
    assert ranges[0][0] == 0 and ranges[-1][1] == L.total
    assert all(a[1] == b[0] for a, b in zip(ranges, ranges[1:]))
    with pytest.raises(ValueError):  # transformer.py:332
        ParamLayout(ModelConfig(vocab_size=8, seq_len=8, dim=64, expand=2.0, n_layers=1, n_heads=1, mlp="gelu"))
    # SURVEY §8f N2 variants: fc1 alone in the gate|up slot, no lm_head parameter when the embeddings are tied
    for mlp, tied in (("mlp", False), ("mlp_relu_sq", True), ("glu", True)):
        cfg_v = ModelConfig(vocab_size=50257, seq_len=64, dim=128, expand=8 / 3, n_layers=2, n_heads=2, mlp=mlp,
                            tie_embeddings=tied)
        Lv = ParamLayout(cfg_v)
        cfg_ov = O.ModelConfig(vocab_size=50257, seq_len=64, dim=128, expand=8 / 3, n_layers=2, n_heads=2, mlp=mlp,
                               tie_embeddings=tied)
        shapes_v = O.param_shapes(cfg_ov)
        assert {v.name: v.shape for v in Lv.views} == shapes_v
        flat_v = torch.zeros(Lv.total)
        for v in Lv.views:
            Lv.param_view(flat_v, v).add_(1.0)
        assert float(flat_v.max()) == 1.0 and int(flat_v.sum()) == sum(int(np.prod(s)) for s in shapes_v.values())
        assert Lv.slots["embed"].rows == (Lv.V_pad if tied else 50257)
    with pytest.raises(NotImplementedError):
        model_config_from_cfg(dict(model="pythia"))
    with pytest.raises(ValueError):
        model_config_from_cfg(dict(model="transformer", vocab_size=8, d_model=64, expand="2", n_layers=1, n_heads=1,
                                   mlp_class="glu", seq_len=8, dtype="int8"))


def test_bucket_groups_partition_the_layers_in_backward_order():
    """Engine's gradient buckets: contiguous layer groups, top layers first, every layer exactly once, at most
    `layers_per_bucket` per group, tapering to single layers at the end of backward."""
    from plaincv_b200.engine import bucket_groups
    assert bucket_groups(24, 8) == [(16, 23), (8, 15), (4, 7), (2, 3), (1, 1), (0, 0)]
    assert bucket_groups(24, 1) == [(i, i) for i in range(23, -1, -1)]
    for n in (1, 2, 3, 4, 12, 24, 25):
        for g in (1, 2, 4, 8, 100):
            groups = bucket_groups(n, g)
            covered = [i for lo, hi in groups for i in range(hi, lo - 1, -1)]
            assert covered == list(range(n - 1, -1, -1)), (n, g, groups)
            assert all(hi - lo + 1 <= g for lo, hi in groups)
            assert groups[-1] == (0, 0)


def test_padded_rows_stay_out_of_the_loss_means():
    """Host side of ragged sequence lengths (model.Transformer._prep_padded / ops._ce_means): with sequences padded from
    T to T_run positions, loss and accuracy are means over the B*T real tokens (_loss_and_acc, train_lm.py:181-186) and
    the pad rows of d(logits) are zeroed; real rows are untouched."""
    from plaincv_b200 import ops
    B, T, T_run, V = 3, 5, 8, 16
    g = torch.Generator().manual_seed(0)
    row_loss = torch.rand(B * T_run, generator=g)
    row_correct = (torch.rand(B * T_run, generator=g) > 0.5).float()
    dlogits = torch.randn(B * T_run, V, generator=g)
    keep = dlogits.clone()
    out2 = ops._ce_means(None, row_loss, row_correct, dlogits, (B, T, T_run))
    real = torch.zeros(B, T_run, dtype=torch.bool)
    real[:, :T] = True
    real = real.view(-1)
    assert torch.allclose(out2[0], row_loss[real].mean()) and torch.allclose(out2[1], row_correct[real].mean())
    assert torch.equal(dlogits[real], keep[real]) and float(dlogits[~real].abs().sum()) == 0.0
    # evaluation (no gradient tensor) takes the same means
    out2e = ops._ce_means(None, row_loss, row_correct, None, (B, T, T_run))
    assert torch.equal(out2e, out2)


def test_product_path_refuses_cpu():
    from plaincv_b200.model import ModelConfig, Transformer
    cfg = ModelConfig(vocab_size=64, seq_len=64, dim=64, expand=8 / 3, n_layers=1, n_heads=1)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        Transformer(cfg, device="cpu")
    from plaincv_b200 import ops
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops._req(torch.zeros(2), torch.float32, "x")


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "plaincv_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, fn)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, re.M), fn


# ------------------------------------------------------------------------------------------------ C ABI
def _declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "plaincv_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(pcv_[a-z0-9_]+)\s*\(", hdr)))


def test_cabi_library_loads_and_exports_every_declared_symbol():
    from plaincv_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    names = _declared_symbols()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/plaincv_b200.h but not exported"
    # the ctypes table binds exactly the declared surface
    assert sorted(_lib.SIGNATURES) == names
    lib.pcv_version.restype = ctypes.c_int
    assert lib.pcv_version() >= 100
    # argument validation happens before any CUDA call: a null pointer is PCV_ERR_ARG, not a crash
    loaded = _lib.load()
    assert loaded.pcv_swiglu_fwd(None, None, 4, 8, None) == -1
    assert loaded.pcv_gemm_bf16(None, None, None, 128, 128, 64, 64, 64, 128, 0, 0, 0, 1, 1.0, None) == -1


def test_sass_uses_tcgen05_and_tma():
    """The GEMM / attention kernels are tcgen05 + TMA (UTC*MMA / UTMALDG in SASS), not mma.sync recompiles."""
    import shutil
    import subprocess
    from plaincv_b200 import _lib
    if shutil.which("cuobjdump") is None or not os.path.exists(_lib.LIB_PATH):
        pytest.skip("cuobjdump or the library is not available")
    sass = subprocess.run(["cuobjdump", "-sass", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "UTCHMMA" in sass and "UTMALDG" in sass and "LDTM" in sass
    assert "HMMA.16816" not in sass


# ------------------------------------------------------------------------------------------------ gloo, world_size 2
def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _ddp_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    sys.path.insert(0, ROOT)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from plaincv_b200.engine import GradReducer
        n = 1000
        ranges = [(0, 300), (300, 650), (650, 1000)]
        g = torch.Generator().manual_seed(100 + rank)
        flat = torch.randn(n, generator=g)
        mine = flat.clone()
        red = GradReducer(flat, ranges)
        # buckets become ready in backward order (last first), one is left for finish() to launch
        red.bucket_ready(2)
        red.bucket_ready(1)
        stat = torch.tensor([float(rank), 10.0 * rank])
        red.reduce_scalars(stat)
        red.finish()
        others = [torch.zeros(n) for _ in range(world)]
        dist.all_gather(others, mine)
        exp = sum(others) / world
        ok = torch.allclose(flat, exp, atol=1e-6) and torch.allclose(stat, torch.tensor([0.5, 5.0]))
        # second round re-uses the reducer (state reset by finish)
        flat.copy_(mine)
        red.finish()
        ok = ok and torch.allclose(flat, exp, atol=1e-6)
        # bf16 on the wire: each rank's bucket is rounded to bf16, the mean is rounded again
        flat.copy_(mine)
        red16 = GradReducer(flat, ranges, comm_dtype=torch.bfloat16)
        red16.bucket_ready(0)
        red16.finish()
        bf = lambda t: t.to(torch.bfloat16).to(torch.float32)  # noqa: E731
        exp16 = bf(sum(bf(o) for o in others) / world)
        ok = ok and torch.equal(flat, exp16) and float((flat - exp).abs().max()) < 2e-2
        # sharding: rank r takes the r-th contiguous slice of the global batch (train_lm.py:151-159)
        glob = torch.arange(8 * 5).view(8, 5)
        from plaincv_b200.batch import shard_for_rank
        sh = shard_for_rank(glob, rank, world)
        parts = [torch.zeros_like(sh) for _ in range(world)]
        dist.all_gather(parts, sh)
        ok = ok and torch.equal(torch.cat(parts), glob)
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


def test_grad_reducer_gloo_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_ddp_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(res) == [(0, True), (1, True)]


def _oracle_ddp_worker(rank, world, port, q):
    """Data-parallel semantics end to end on CPU: each rank runs the ORACLE on its shard, the product's GradReducer
    averages the flat gradients over gloo; the result must equal the oracle on the whole batch (lax.pmean)."""
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    sys.path.insert(0, ROOT)
    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from plaincv_b200.engine import GradReducer
        cfg = O.ModelConfig(vocab_size=211, seq_len=32, dim=64, expand=8 / 3, n_layers=2, n_heads=1)
        params = O.init_params(cfg, seed=0)
        batch = O.synthetic_batch(4, 32, 211, seed=3)
        inputs, labels, _ = O.prepare_batch(batch, 32, False)
        from plaincv_b200.batch import shard_for_rank
        g_loc, loss, acc = O.compute_grads(params, cfg, shard_for_rank(inputs, rank, world),
                                           shard_for_rank(labels, rank, world))
        names = sorted(g_loc)
        flat = torch.cat([g_loc[k].reshape(-1) for k in names])
        half = flat.numel() // 2
        red = GradReducer(flat, [(0, half), (half, flat.numel())])
        red.bucket_ready(1)
        stat = torch.stack([loss, acc]).float()
        red.reduce_scalars(stat)
        red.finish()
        g_all, loss_all, _ = O.compute_grads(params, cfg, inputs, labels)
        flat_all = torch.cat([g_all[k].reshape(-1) for k in names])
        ok = torch.allclose(flat, flat_all, rtol=1e-4, atol=1e-7) and abs(float(stat[0]) - float(loss_all)) < 1e-5
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


def test_data_parallel_mean_equals_whole_batch_gloo_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_oracle_ddp_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(res) == [(0, True), (1, True)]
