"""Synthetic parquet corpus (BASELINE.json: "synthetic parquet token data of the named
shape").  One `text` column, like the reference's training file (utils.py:114-119).
Documents are sequences of `t<int>` words, which the offline tokenizer maps to exact
token ids; a Zipf-like skew gives the loss something to learn."""
from __future__ import annotations

import os

import numpy as np
import pyarrow as pa
import pyarrow.parquet as pq


def write_synthetic_parquet(path: str, num_docs: int = 256, min_len: int = 64, max_len: int = 4096,
                            vocab_size: int = 131072, seed: int = 1234) -> str:
    rng = np.random.default_rng(seed)
    docs = []
    for _ in range(num_docs):
        n = int(rng.integers(min_len, max_len + 1))
        # mixture: frequent small ids + a uniform tail over the whole vocabulary
        head = rng.zipf(1.3, size=n) % 4096
        tail = rng.integers(0, vocab_size - 4, size=n)
        ids = np.where(rng.random(n) < 0.7, head, tail)
        docs.append(" ".join(f"t{int(i)}" for i in ids))
    os.makedirs(os.path.dirname(os.path.abspath(path)) or ".", exist_ok=True)
    tmp = path + ".tmp"
    pq.write_table(pa.table({"text": docs}), tmp)
    os.replace(tmp, path)
    return path
