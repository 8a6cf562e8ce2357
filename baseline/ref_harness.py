"""Reference arm of bench.py: runs the UNMODIFIED reference (baseline/_ref, byte-identical to
/root/reference/{train,utils,model,dataset}.py) through its own public entry point `train.train(args)`.

Nothing of the new framework is on this path: no model, kernel, engine, dataset or tokenizer of
`fault_tolerant_llm_training_b200` is imported.  What the harness adds is only what the reference needs
to run offline on this box:

  * a stand-in for `AutoTokenizer.from_pretrained` (no HF hub here): a word-hash tokenizer with the
    Mistral-Nemo vocabulary size and the `encode_plus` method the reference calls (transformers 5
    removed it);
  * a synthetic parquet with a `text` column;
  * timing: the reference's `DataLoader` name is replaced by a subclass whose iterator notes when batch
    number W is requested (start of the timed region; CUDA event + wall clock) — the loop body itself
    is the reference's, untouched; the end of the region is the SystemExit that train() raises after
    "Training completed" (train.py:118-119);
  * N > 1 (the reference has no multi-GPU path): the model class name seen by train.py is a subclass
    that routes forward() through torch DistributedDataParallel — "the same model wrapped in torch DDP
    over NCCL" of BASELINE.md §3.  Every rank runs the reference loop on the same data.
"""
from __future__ import annotations

import argparse
import os
import re
import sys
import time
import zlib

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")


class ShimTokenizer:
    """Offline stand-in with the interface the reference uses (dataset.py:29-35, train.py:28-30, 51)."""

    vocab_size = 131072
    pad_token_id = 3
    bos_token_id = 1
    _word = re.compile(r"\S+")

    def _id(self, w: str) -> int:
        if len(w) > 1 and w[0] == "t" and w[1:].isdigit():
            return 4 + int(w[1:]) % (self.vocab_size - 4)
        return 4 + zlib.crc32(w.encode()) % (self.vocab_size - 4)

    def encode_plus(self, text, max_length=None, padding=False, truncation=False, padding_side="right", **_):
        ids = [self.bos_token_id] + [self._id(w) for w in self._word.findall(text)]
        if truncation and max_length:
            ids = ids[:max_length]
        if padding == "max_length" and max_length and len(ids) < max_length:
            ids = ids + [self.pad_token_id] * (max_length - len(ids))
        return {"input_ids": ids, "attention_mask": [1] * len(ids)}

    __call__ = encode_plus


def write_parquet(path: str, docs: int, words: int, seed: int = 1234) -> None:
    import numpy as np
    import pyarrow as pa
    import pyarrow.parquet as pq

    rng = np.random.default_rng(seed)
    rows = []
    for _ in range(docs):
        head = rng.zipf(1.3, size=words) % 4096
        tail = rng.integers(0, 131068, size=words)
        ids = np.where(rng.random(words) < 0.7, head, tail)
        rows.append(" ".join(f"t{int(i)}" for i in ids))
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    pq.write_table(pa.table({"text": rows}), path + ".tmp")
    os.replace(path + ".tmp", path)


def reference_available() -> str:
    """'' when the reference install is usable, else the reason."""
    for f in ("train.py", "utils.py", "model.py", "dataset.py"):
        if not os.path.exists(os.path.join(REF_DIR, f)):
            return f"baseline/_ref/{f} missing (pip install --target baseline/_ref not done)"
    return ""


def run_reference(n_gpus: int, steps: int, warmup: int, seq_len: int, batch: int, logging_frequency: int,
                  data_path: str, clock_sampler=None) -> dict:
    """One call of the reference's train(args) for warmup+steps steps; returns timings of the last `steps`."""
    import torch

    if REF_DIR not in sys.path:
        sys.path.insert(0, REF_DIR)
    os.environ.setdefault("WORKDIR", "/tmp/ftl_ref_workdir")
    import transformers

    transformers.AutoTokenizer.from_pretrained = staticmethod(lambda *a, **k: ShimTokenizer())
    import train as ref_train  # the reference's train.py

    ref_train.AutoTokenizer = transformers.AutoTokenizer
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local_rank)

    if world > 1:
        import torch.distributed as dist
        from torch.nn.parallel import DistributedDataParallel as DDP

        if not dist.is_initialized():
            dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local_rank}"))
        RefTransformer = ref_train.Transformer

        class _Body(torch.nn.Module):
            def __init__(self, outer):
                super().__init__()
                self.outer = [outer]
                for n, p in outer.named_parameters():
                    self.register_parameter(n.replace(".", "__"), p)

            def forward(self, tokens):
                return RefTransformer.forward(self.outer[0], tokens)

        class DDPTransformer(RefTransformer):
            def forward(self, tokens):
                ddp = self.__dict__.get("_ddp")
                if ddp is None:
                    ddp = DDP(_Body(self), device_ids=[local_rank], gradient_as_bucket_view=True)
                    self.__dict__["_ddp"] = ddp
                return ddp(tokens)

        ref_train.Transformer = DDPTransformer

    state = {"calls": 0, "t0": None, "start": torch.cuda.Event(enable_timing=True),
             "end": torch.cuda.Event(enable_timing=True)}
    BaseLoader = ref_train.DataLoader

    class TimedLoader(BaseLoader):
        def __iter__(self):
            inner = super().__iter__()

            class _It:
                def __iter__(s):
                    return s

                def __next__(s):
                    if state["calls"] == warmup:
                        if world > 1:
                            torch.distributed.barrier()
                        torch.cuda.synchronize()
                        if clock_sampler is not None:
                            clock_sampler.start()
                        state["t0"] = time.perf_counter()
                        state["start"].record()
                    state["calls"] += 1
                    return next(inner)

            return _It()

    ref_train.DataLoader = TimedLoader
    args = argparse.Namespace(
        dataset=data_path, checkpoint_path="/tmp/ftl_ref_workdir/checkpoints", checkpoint_id="",
        tokenizer_name_or_path="offline-shim", sequence_length=seq_len, batch_size=batch, fused_optimizer=False,
        learning_rate=5e-5, lr_warmup_steps=100, training_steps=warmup + steps, logging_frequency=logging_frequency,
        grad_max_norm=1, model_dtype="bf16", compile=False, raise_error=False, error_step=10 ** 9)
    ref_train.init_logger() if hasattr(ref_train, "init_logger") and rank == 0 else None
    t_build = time.perf_counter()
    try:
        ref_train.train(args)
    except SystemExit:
        pass
    state["end"].record()
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    if clock_sampler is not None:
        clock_sampler.stop()
    if world > 1:
        torch.distributed.barrier()
    if state["t0"] is None or state["calls"] < warmup + steps:
        raise RuntimeError(f"reference loop ended early after {state['calls']} batches")
    return {"device_ms": state["start"].elapsed_time(state["end"]), "wall_ms": (t1 - state["t0"]) * 1e3,
            "total_s": t1 - t_build}
