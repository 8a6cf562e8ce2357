"""Datasets: same three public classes as the reference (dataset.py:10-101), with the
same constructor signatures and the same `(inputs, labels)` contract, plus O(1)
resumable position (`state_dict` / `load_state_dict`) so that resume does not
re-tokenise every skipped batch the way train.py:36-39 does (6.9 s for 427
batches, logs/output_444671.out:7-8).
"""
from __future__ import annotations

import queue
import threading
from dataclasses import dataclass
from typing import Dict, List, Optional, Tuple

import pyarrow.parquet as pq
import torch
from torch.utils.data import Dataset, IterableDataset


def _tokenize(tokenizer, text: str, **kw) -> Dict[str, List[int]]:
    # transformers>=5 removed encode_plus; __call__ has the same keywords.
    return tokenizer(text, **kw)


class ParquetDataset(Dataset):
    """Map-style, one document per sample, right-padded/truncated to
    sequence_length+1 tokens; index wraps modulo the file length
    (reference: dataset.py:10-35)."""

    def __init__(self, parquet_file: str, tokenizer, sequence_length: int, training_samples: int):
        self.parquet_ds = pq.read_table(parquet_file, memory_map=True)
        self.real_length = len(self.parquet_ds)
        if self.real_length == 0:
            raise ValueError(f"{parquet_file} has no rows")
        self._text = self.parquet_ds["text"]
        self.tokenizer = tokenizer
        self.sequence_length = sequence_length
        self.training_samples = training_samples

    def __len__(self) -> int:
        return self.training_samples

    def __getitem__(self, idx: int) -> Dict[str, List[int]]:
        sample_str = str(self._text[idx % self.real_length])
        return _tokenize(
            self.tokenizer, sample_str,
            max_length=self.sequence_length + 1,
            padding="max_length", truncation=True, padding_side="right",
        )


@dataclass
class CollatorForCLM:
    """Shift-by-one causal-LM collation with pad positions masked to -100
    (reference: dataset.py:38-53).  Returns a tuple (inputs, labels), both int64
    of shape [b, sequence_length]."""

    sequence_length: int
    pad_token_id: int

    def __call__(self, examples: List[Dict[str, List[int]]]) -> Tuple[torch.Tensor, torch.Tensor]:
        ids = torch.tensor([ex["input_ids"] for ex in examples], dtype=torch.long)  # (b, s+1)
        if ids.shape[1] != self.sequence_length + 1:
            raise ValueError(f"expected {self.sequence_length + 1} tokens per example, got {ids.shape[1]}")
        inputs = ids[:, :-1].clone()
        labels = ids[:, 1:].clone()
        labels[labels == self.pad_token_id] = -100
        return inputs, labels


class IterableParquetDataset(IterableDataset):
    """Infinite packing dataset (reference: dataset.py:56-101).

    Documents are tokenised and concatenated until sequence_length+1 tokens are
    available; labels are masked at and after every BOS exactly as
    dataset.py:99-100 does.  Differences from the reference, all from SURVEY Q6:
    the cursor always advances (a document of >= sequence_length+1 tokens stalls
    the reference forever), tokens left over after cutting a sample are carried
    into the next sample instead of being dropped and re-read, `__iter__` keeps
    the position, and the position is checkpointable.  `stride`/`offset` give
    each data-parallel rank its own document stream.
    """

    def __init__(self, parquet_file: str, tokenizer, sequence_length: int, bos_token_id: int = 1,
                 offset: int = 0, stride: int = 1):
        self.parquet_ds = pq.read_table(parquet_file, memory_map=True)
        self.real_length = len(self.parquet_ds)
        if self.real_length == 0:
            raise ValueError(f"{parquet_file} has no rows")
        self._text = self.parquet_ds["text"]
        self.tokenizer = tokenizer
        self.sequence_length = sequence_length
        self.bos_token_id = bos_token_id
        self.offset, self.stride = offset, stride
        self.current_index = 0
        self.token_buffer: List[int] = []

    def __iter__(self):
        return self

    def __next__(self) -> Tuple[torch.Tensor, torch.Tensor]:
        need = self.sequence_length + 1
        while len(self.token_buffer) < need:
            doc = (self.offset + self.current_index * self.stride) % self.real_length
            toks = _tokenize(self.tokenizer, str(self._text[doc]), padding=False,
                             truncation=True, max_length=need)
            self.token_buffer.extend(toks["input_ids"])
            self.current_index += 1
        window, self.token_buffer = self.token_buffer[:need], self.token_buffer[need:]
        inputs = torch.tensor(window[:-1], dtype=torch.long)
        labels = torch.tensor(window[1:], dtype=torch.long)
        labels[inputs == self.bos_token_id] = -100
        labels[labels == self.bos_token_id] = -100
        return inputs, labels

    def state_dict(self) -> dict:
        return {"current_index": self.current_index, "token_buffer": list(self.token_buffer),
                "offset": self.offset, "stride": self.stride}

    def load_state_dict(self, state: dict) -> None:
        self.current_index = int(state["current_index"])
        self.token_buffer = list(state["token_buffer"])


class BatchStream:
    """Deterministic, resumable batch source for the training loop.

    padded mode: batch `i` of rank `r` is samples
    [i*B*W + r*B, i*B*W + (r+1)*B) of the ParquetDataset, i.e. W ranks at batch B
    consume exactly the samples one rank at batch B*W would.
    Restoring is O(1): set `batches_consumed`.

    packed mode: wraps an IterableParquetDataset whose documents are strided by
    rank; its own state_dict is stored.

    A background thread tokenises ahead (the reference tokenises on the training
    thread); batches are delivered in pinned memory when asked.  Pinned
    batches are views into a ring of `prefetch + 8` page-locked buffers allocated
    once: a consumer that wants to keep a batch beyond the next few must copy it
    (the trainers copy every batch into their own staging memory inside `step()`).
    """

    def __init__(self, dataset, batch_size: int, collator: Optional[CollatorForCLM] = None,
                 rank: int = 0, world_size: int = 1, pin_memory: bool = False, prefetch: int = 2):
        self.dataset = dataset
        self.collator = collator
        self.batch_size = batch_size
        self.rank, self.world_size = rank, world_size
        self.pin_memory = pin_memory and torch.cuda.is_available()
        self.packed = isinstance(dataset, IterableParquetDataset)
        self.batches_consumed = 0
        self._prefetch = 0 if self.packed else prefetch
        self._q: Optional[queue.Queue] = None
        self._thread: Optional[threading.Thread] = None
        self._stop = threading.Event()
        self._history = {}   # packed mode: dataset state before batch i, for the last few batches (rewind)

    # -- production ---------------------------------------------------------
    def _make(self, batch_index: int) -> Tuple[torch.Tensor, torch.Tensor]:
        if self.packed:
            self._history[batch_index] = self.dataset.state_dict()
            for k in [k for k in self._history if k < batch_index - 8]:
                del self._history[k]
            pairs = [next(self.dataset) for _ in range(self.batch_size)]
            inputs = torch.stack([p[0] for p in pairs])
            labels = torch.stack([p[1] for p in pairs])
        else:
            g = self.batch_size * self.world_size
            first = batch_index * g + self.rank * self.batch_size
            examples = [self.dataset[first + j] for j in range(self.batch_size)]
            inputs, labels = self.collator(examples)
        if self.pin_memory:
            inputs, labels = self._to_pinned(inputs, labels)
        return inputs, labels

    def _to_pinned(self, inputs: torch.Tensor, labels: torch.Tensor):
        """Copy into a small ring of page-locked buffers allocated once.  `tensor.pin_memory()` per batch is a
        cudaHostAlloc on the producer thread every step: slow, and it takes driver locks the training thread needs for
        its ~800 kernel launches.  The ring is deeper than the prefetch queue plus the batches a consumer can still be
        reading (the trainers copy a batch out before they ask for the next one)."""
        key = (tuple(inputs.shape), inputs.dtype, tuple(labels.shape), labels.dtype)
        if getattr(self, "_pin_key", None) != key:
            n = self._prefetch + 8
            self._pin_ring = [(torch.empty(inputs.shape, dtype=inputs.dtype).pin_memory(),
                               torch.empty(labels.shape, dtype=labels.dtype).pin_memory()) for _ in range(n)]
            self._pin_key, self._pin_next = key, 0
        a, b = self._pin_ring[self._pin_next]
        self._pin_next = (self._pin_next + 1) % len(self._pin_ring)
        a.copy_(inputs)
        b.copy_(labels)
        return a, b

    def _worker(self, start: int) -> None:
        i = start
        while not self._stop.is_set():
            try:
                item = (i, self._make(i))
            except Exception as e:  # surfaced on the consumer side
                item = (i, e)
            while not self._stop.is_set():
                try:
                    self._q.put(item, timeout=0.1)
                    break
                except queue.Full:
                    continue
            i += 1

    def _start(self) -> None:
        # The producer tokenises in Python; with the default 5 ms GIL switch interval it can hold the interpreter for
        # 5 ms at a time while the training thread is in the middle of its per-step launch sequence.
        import sys
        if sys.getswitchinterval() > 5e-4:
            sys.setswitchinterval(5e-4)
        self._stop.clear()
        self._q = queue.Queue(maxsize=self._prefetch)
        self._thread = threading.Thread(target=self._worker, args=(self.batches_consumed,), daemon=True)
        self._thread.start()

    def close(self) -> None:
        if self._thread is not None:
            self._stop.set()
            self._thread.join(timeout=5)
            self._thread, self._q = None, None

    # -- consumption --------------------------------------------------------
    def __iter__(self):
        return self

    def __next__(self) -> Tuple[torch.Tensor, torch.Tensor]:
        if self._prefetch <= 0:
            batch = self._make(self.batches_consumed)  # packed mode: index only labels the history entry
        else:
            if self._thread is None:
                self._start()
            idx, batch = self._q.get()
            if isinstance(batch, Exception):
                raise batch
            assert idx == self.batches_consumed, (idx, self.batches_consumed)
        self.batches_consumed += 1
        return batch

    # -- checkpointing ------------------------------------------------------
    def state_dict(self) -> dict:
        st = {"kind": "packed" if self.packed else "padded",
              "batches_consumed": self.batches_consumed,
              "batch_size": self.batch_size, "world_size": self.world_size, "rank": self.rank}
        if self.packed:
            st["dataset"] = self.dataset.state_dict()
        return st

    def load_state_dict(self, state: dict) -> None:
        self.close()
        self.batches_consumed = int(state["batches_consumed"])
        if self.packed and "dataset" in state:
            self.dataset.load_state_dict(state["dataset"])

    def restore(self, saved) -> None:
        """Restore from what a checkpoint holds: one state (single process) or the list of all ranks' states.

        Same world size and batch size: this rank's own state.  A different geometry is accepted only where the
        position can be converted exactly — padded mode, where batch i of rank r is a fixed slice of the global
        sample order: samples consumed = batches * B_old * W_old must be a whole number of new global batches.
        Packed mode strides documents by rank and carries a per-rank token buffer, so it cannot be re-cut:
        refused with a clear error rather than silently skipping or repeating samples."""
        states = saved if isinstance(saved, list) else [saved]
        w_old = int(states[0].get("world_size", len(states)))
        b_old = int(states[0].get("batch_size", self.batch_size))
        if w_old == self.world_size and b_old == self.batch_size and len(states) > (self.rank if len(states) > 1 else 0):
            self.load_state_dict(states[self.rank] if len(states) > 1 else states[0])
            return
        kind = states[0].get("kind", "padded")
        if self.packed or kind == "packed":
            raise ValueError(f"checkpoint was written by {w_old} rank(s) at batch size {b_old} with the packed dataset; "
                             f"its position cannot be converted to {self.world_size} rank(s) at batch size "
                             f"{self.batch_size} (resume with the original geometry or use --dataset-type padded)")
        consumed = {int(st["batches_consumed"]) for st in states}
        if len(consumed) != 1:
            raise ValueError(f"ranks of the checkpoint disagree on the data position: {sorted(consumed)}")
        samples = consumed.pop() * b_old * w_old
        g_new = self.batch_size * self.world_size
        if samples % g_new:
            raise ValueError(f"checkpoint consumed {samples} samples ({w_old} rank(s) x batch {b_old}); that is not a "
                             f"whole number of global batches of {g_new} ({self.world_size} rank(s) x batch {self.batch_size})")
        self.close()
        self.batches_consumed = samples // g_new

    def rewind(self, batches_consumed: int) -> None:
        """Move the position back to just before batch `batches_consumed` (a step that was fetched but whose
        update never committed).  Padded mode is O(1); packed mode restores the recorded dataset state."""
        if batches_consumed == self.batches_consumed:
            return
        self.close()
        if self.packed:
            if batches_consumed not in self._history:
                raise ValueError(f"no recorded dataset state for batch {batches_consumed}")
            self.dataset.load_state_dict(self._history[batches_consumed])
        self.batches_consumed = batches_consumed

    def fast_forward(self, batches: int) -> None:
        """Position restore for checkpoints that carry no dataset state (files
        written by the reference): padded mode is O(1); packed mode replays."""
        self.close()
        if self.packed:
            for _ in range(batches):
                self._make(0)
        self.batches_consumed = batches


def _smoke(argv=None) -> None:
    """Dataset smoke run (reference: `python dataset.py`, dataset.py:104-166): builds both datasets, prints the
    batch shapes and the share of label positions masked out of the loss.  Works offline: without --dataset a
    synthetic parquet is written, without a reachable HF tokenizer the offline tokenizer is used."""
    import argparse
    import os
    import tempfile

    from .synthetic import write_synthetic_parquet
    from .tokenizer import load_tokenizer

    ap = argparse.ArgumentParser()
    ap.add_argument("--dataset", default="")
    ap.add_argument("--tokenizer-name-or-path", default="offline")
    ap.add_argument("--sequence-length", type=int, default=4096)
    ap.add_argument("--batch-size", type=int, default=32)
    a = ap.parse_args(argv)
    tok = load_tokenizer(a.tokenizer_name_or_path)
    path = a.dataset or write_synthetic_parquet(os.path.join(tempfile.mkdtemp(), "smoke.parquet"), num_docs=64,
                                                max_len=2 * a.sequence_length, vocab_size=tok.vocab_size)
    ds = ParquetDataset(path, tok, a.sequence_length, training_samples=a.batch_size)
    print("Decoded sample:", tok.decode(ds[0]["input_ids"][:20]))
    padded = BatchStream(ds, a.batch_size, CollatorForCLM(a.sequence_length, tok.pad_token_id), prefetch=0)
    packed = BatchStream(IterableParquetDataset(path, tok, a.sequence_length, bos_token_id=tok.bos_token_id or 1),
                         a.batch_size)
    for name, stream in (("ParquetDataset + CollatorForCLM", padded), ("IterableParquetDataset", packed)):
        x, y = next(stream)
        ignored = int((y == -100).sum())
        print(f"{name}: input {tuple(x.shape)} labels {tuple(y.shape)} | ignored tokens in loss: {ignored} of {y.numel()} "
              f"({ignored / y.numel() * 100:.2f}%)")


if __name__ == "__main__":
    _smoke()
