#!/usr/bin/env python
"""Writes tests/golden/hf_tokens/{plain,docs}: two tiny datasets in the on-disk format the reference trains from
(`datasets.Dataset.save_to_disk`, read back by data/lm_loader.py:17 with load_from_disk), produced with the real
`datasets` library so that plaincv_b200.feed.ArrowTokenDataset is tested against genuine files.
Row i of `plain`: input_ids[j] = (i * 131 + j * 7) % 50257, 33 tokens (seq_len 32 + 1).
`docs` adds docs_lengths: row i is cut into documents of lengths cycling through (5, 9, 3, 16) from offset i,
truncated so that they sum to 33.   usage: python tools/make_golden_dataset.py"""
import os
import shutil

import numpy as np
from datasets import Dataset

ROOT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "hf_tokens")
N, W = 23, 33


def row_ids(i):
    return ((i * 131 + np.arange(W) * 7) % 50257).astype(np.int32)


def row_docs(i):
    lens, k, tot = [], i, 0
    cyc = (5, 9, 3, 16)
    while tot < W:
        n = min(cyc[k % 4], W - tot)
        lens.append(n)
        tot += n
        k += 1
    return lens


def main():
    shutil.rmtree(ROOT, ignore_errors=True)
    os.makedirs(ROOT)
    ids = [row_ids(i).tolist() for i in range(N)]
    Dataset.from_dict({"input_ids": ids}).save_to_disk(os.path.join(ROOT, "plain"))
    Dataset.from_dict({"input_ids": ids, "docs_lengths": [row_docs(i) for i in range(N)]}).save_to_disk(
        os.path.join(ROOT, "docs"), num_shards=2)
    for d in ("plain", "docs"):
        print(d, sorted(os.listdir(os.path.join(ROOT, d))))


if __name__ == "__main__":
    main()
