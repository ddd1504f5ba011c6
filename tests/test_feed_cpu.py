"""Batch feeding on CPU: plaincv_b200.feed against the committed datasets under
tests/golden/hf_tokens (written by tools/make_golden_dataset.py with the real `datasets` library, i.e. the on-disk
format data/lm_loader.py:17 reads), against re-statements of the reference's samplers (data/lm_datasampler.py) and of
its loader loop (lm_loader.py:30-50, train_lm.py:356-380)."""
import os

import numpy as np
import pytest
import torch

from plaincv_b200 import batch as PB
from plaincv_b200 import feed

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "hf_tokens")
N, W = 23, 33


def _row_ids(i):
    return ((i * 131 + np.arange(W) * 7) % 50257).astype(np.int64)


def _row_docs(i):
    lens, k, tot = [], i, 0
    cyc = (5, 9, 3, 16)
    while tot < W:
        n = min(cyc[k % 4], W - tot)
        lens.append(n)
        tot += n
        k += 1
    return lens


def test_arrow_dataset_reads_save_to_disk_format():
    ds = feed.ArrowTokenDataset(os.path.join(GOLD, "plain"))
    assert len(ds) == N and ds.column_names == ["input_ids"]
    for i in (0, 7, N - 1, -1):
        assert np.array_equal(np.asarray(ds[i]["input_ids"]), _row_ids(i % N))
    dd = feed.ArrowTokenDataset(os.path.join(GOLD, "docs"))  # two arrow shards on disk
    assert len(dd) == N and set(dd.column_names) == {"input_ids", "docs_lengths"}
    for i in range(N):
        r = dd[i]
        assert np.array_equal(np.asarray(r["input_ids"]), _row_ids(i)) and r["docs_lengths"] == _row_docs(i)
    b = dd.rows([3, 20, 11])
    assert b["input_ids"].shape == (3, W) and b["docs_lengths"] == [_row_docs(3), _row_docs(20), _row_docs(11)]
    with pytest.raises(ValueError):
        feed.ArrowTokenDataset(os.path.join(GOLD, "nowhere"))
    with pytest.raises(IndexError):
        ds[N]


def test_arrow_dataset_matches_datasets_library_when_present():
    datasets = pytest.importorskip("datasets")
    ref = datasets.load_from_disk(os.path.join(GOLD, "docs"))
    ds = feed.ArrowTokenDataset(os.path.join(GOLD, "docs"))
    assert len(ref) == len(ds)
    for i in range(len(ref)):
        assert list(ref[i]["input_ids"]) == np.asarray(ds[i]["input_ids"]).tolist()
        assert list(ref[i]["docs_lengths"]) == ds[i]["docs_lengths"]
    for shards in (2, 3, 5):  # lm_loader.py:23-28: one contiguous shard per process
        for k in range(shards):
            a, b = ref.shard(num_shards=shards, index=k, contiguous=True), ds.shard(shards, k)
            assert len(a) == len(b)
            if len(a):
                assert list(a[0]["input_ids"]) == np.asarray(b[0]["input_ids"]).tolist()
                assert list(a[len(a) - 1]["input_ids"]) == np.asarray(b[-1]["input_ids"]).tolist()
    assert len(ds.take(5)) == 5 and len(ds.take(100)) == N


def test_index_streams_follow_the_reference_samplers():
    n, B = 103, 8
    # StatefulSequentialSampler (lm_datasampler.py:19-31)
    assert feed.epoch_indices(n, "stateful_sequential", batch_size=B, start_idx=3).tolist() == list(range(24, n))
    assert feed.epoch_indices(n, "sequential").tolist() == list(range(n))
    # StatefulRandomSampler (:34-70): one persistent Generator, a fresh permutation per epoch, offset on the first
    rng_ref = np.random.default_rng(17)
    e0 = rng_ref.permutation(np.arange(n, dtype=np.int64))[2 * B:]
    e1 = rng_ref.permutation(np.arange(n, dtype=np.int64))
    rng = np.random.default_rng(17)
    assert np.array_equal(feed.epoch_indices(n, "stateful_random", batch_size=B, start_idx=2, rng=rng), e0)
    assert np.array_equal(feed.epoch_indices(n, "stateful_random", batch_size=B, rng=rng), e1)
    with pytest.raises(ValueError):
        feed.epoch_indices(n, "stateful_random")
    with pytest.raises(NotImplementedError):
        feed.epoch_indices(n, "weighted")
    # StatefulJaxDistributedSampler (:73-146)
    for rank in range(4):
        got = feed.distributed_indices(n, rank, 4, B, seed=5, epoch=2, start_iter=1)
        num = (n - n % 4) // 4
        base = np.arange(rank * num, (rank + 1) * num, dtype=np.int64)
        exp = base[np.random.default_rng(7).permutation(num)][B:]
        assert np.array_equal(got, exp)
    allr = np.concatenate([feed.distributed_indices(n, r, 4, B, shuffle_within_rank=False) for r in range(4)])
    assert allr.tolist() == list(range(100))  # drop_last: 103 -> 100, disjoint contiguous blocks
    padded = feed.distributed_indices(n, 3, 4, B, shuffle_within_rank=False, drop_last=False)
    assert len(padded) == 26 and padded[-1] == (4 * 26 - 1) % n


@pytest.mark.parametrize("use_doc", [False, True])
def test_feeder_yields_the_loader_loop(use_doc):
    """BatchFeeder == DataLoader(drop_last=True) + _next_batch wrap-around + _prepare_batch, staged ahead."""
    ds = feed.ArrowTokenDataset(os.path.join(GOLD, "docs" if use_doc else "plain"))
    B, T = 4, 32
    f = feed.BatchFeeder(ds, B, T, use_doc_mask=use_doc, depth=2)
    per_epoch = N // B
    try:
        for k in range(2 * per_epoch + 2):  # runs through two wrap-arounds
            item = next(f)
            rows = [(k % per_epoch) * B + j for j in range(B)]
            ids = np.stack([_row_ids(i) for i in rows])
            exp = {"input_ids": ids}
            if use_doc:
                exp["docs_lengths"] = [_row_docs(i) for i in rows]
            inputs, labels, ds_exp = PB.prepare_batch(exp, T, use_doc)
            assert torch.equal(item.inputs, inputs) and torch.equal(item.labels, labels)
            if use_doc:
                assert torch.equal(item.doc_start, ds_exp)
            else:
                assert item.doc_start is None
            f.release(item)
        st = f.state()
        assert st["micro_batches"] == 2 * per_epoch + 2 and st["epoch"] == 2 and st["start_idx"] == 2
    finally:
        f.close()


@pytest.mark.parametrize("reapply", [True, False])
def test_feeder_resumes_and_shuffles_like_the_stateful_sampler(reapply):
    """Resume at start_idx = 2 with the persistent-Generator shuffle of StatefulRandomSampler
    (lm_datasampler.py:33-66).  reapply_start_idx=True (default) is the reference's behaviour: `__iter__` slices
    `indices[self.start:]` on EVERY epoch (:59-63); False applies the offset to the first epoch only."""
    ds = feed.ArrowTokenDataset(os.path.join(GOLD, "plain"))
    B, T = 3, 32
    f = feed.BatchFeeder(ds, B, T, sampler="stateful_random", sampler_seed=9, start_idx=2, reapply_start_idx=reapply)
    try:
        rng = np.random.default_rng(9)
        order = rng.permutation(np.arange(N, dtype=np.int64))[2 * B:]
        nxt = rng.permutation(np.arange(N, dtype=np.int64))
        if reapply:
            nxt = nxt[2 * B:]
        seq = [order[k * B:(k + 1) * B] for k in range(len(order) // B)] + [nxt[:B]]
        for rows in seq:
            item = next(f)
            assert torch.equal(item.inputs, torch.from_numpy(np.stack([_row_ids(i) for i in rows])[:, :-1].astype(np.int32)))
            f.release(item)
        st = f.state()
        assert st["micro_batches"] == 2 + len(seq) and st["epoch"] == 1
        assert st["start_idx"] == (3 if reapply else 1)  # one batch into the second epoch
    finally:
        f.close()
    with pytest.raises(ValueError):
        feed.BatchFeeder(ds, 64, T)


def test_token_ids_outside_the_vocabulary_are_rejected_on_the_host():
    """ids index the embedding table / logits unchecked in the kernels, so the host contract rejects them."""
    from plaincv_b200.batch import prepare_batch
    ids = np.arange(2 * 9, dtype=np.int32).reshape(2, 9)
    prepare_batch({"input_ids": ids}, 8, False, vocab_size=18)
    with pytest.raises(ValueError, match="token ids must lie in"):
        prepare_batch({"input_ids": ids}, 8, False, vocab_size=17)
    neg = ids.copy()
    neg[0, 0] = -1
    with pytest.raises(ValueError, match="token ids must lie in"):
        prepare_batch({"input_ids": neg}, 8, False, vocab_size=100)
    ds = feed.InMemoryTokenDataset(np.full((4, 9), 50, dtype=np.int32))
    f = feed.BatchFeeder(ds, 2, 8, vocab_size=50, depth=1)
    try:
        with pytest.raises(RuntimeError, match="producer failed"):
            next(f)
    finally:
        f.close()


def test_in_memory_dataset_feeds_like_the_arrow_one():
    """InMemoryTokenDataset (synthetic-data mode): same row / collate contract as ArrowTokenDataset, and BatchFeeder
    over it yields exactly prepare_batch of consecutive row blocks, wrapping around at the end (drop_last)."""
    import torch
    from plaincv_b200.batch import prepare_batch
    from plaincv_b200.feed import BatchFeeder, InMemoryTokenDataset
    ds = InMemoryTokenDataset.synthetic(7, 64, 100, seed=3, doc_mean=20)
    assert len(ds) == 7 and ds[0]["input_ids"].shape == (65,) and sum(ds[-1]["docs_lengths"]) == 65
    r = ds.rows([4, 1])
    assert r["input_ids"].shape == (2, 65) and r["docs_lengths"] == [ds[4]["docs_lengths"], ds[1]["docs_lengths"]]
    with pytest.raises(IndexError):
        ds[7]
    with pytest.raises(ValueError):
        InMemoryTokenDataset(np.zeros((3, 65), dtype=np.int32), docs_lengths=[[65]])
    f = BatchFeeder(ds, 2, 64, use_doc_mask=True, device=None, depth=2)
    try:
        for blk in ([0, 1], [2, 3], [4, 5], [0, 1]):  # 7 rows, batches of 2: row 6 is dropped, then the stream wraps
            item = next(f)
            ref = prepare_batch(ds.rows(blk), 64, True)
            assert torch.equal(item.inputs, ref[0]) and torch.equal(item.labels, ref[1])
            assert torch.equal(item.doc_start, ref[2])
            f.release(item)
    finally:
        f.close()
