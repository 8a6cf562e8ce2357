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
    number W is requested (start of the device-timed region), when batch W+K is requested (end of it; the
    harness then switches `args.logging_frequency` to 1 so that the next K steps read their loss on the host
    every step: the end-to-end region) — the loop body itself is the reference's, untouched; the end of the
    second region is the SystemExit that train() raises after "Training completed" (train.py:118-119).
    One call of train() therefore yields both numbers and the 8 B model is built once per arm;
  * N > 1 (the reference has no multi-GPU path): the model class name seen by train.py is a subclass of the
    reference's `Transformer` that
      - "ddp":  routes forward() through torch DistributedDataParallel, or
      - "fsdp": applies `torch.distributed.fsdp.fully_shard` per TransformerBlock and to the root when
                train.py moves the model to the device (train.py:59), before its optimizer is built
    — "the same model wrapped in torch DDP/FSDP over NCCL" of BASELINE.md §3.

The names patched on the reference's `train` module are captured ONCE in their pristine form and put back after
every call, so `run_reference` can be called any number of times in one process (round 1 wrapped the wrapper on
the second call and recursed).
"""
from __future__ import annotations

import argparse
import contextlib
import gc
import os
import re
import shutil
import subprocess
import sys
import tempfile
import time
import zlib

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
REF_FILES = ("train.py", "utils.py", "model.py", "dataset.py")
REF_SOURCE = "/root/reference"


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


_PYPROJECT = """[build-system]
requires = ["setuptools"]
build-backend = "setuptools.build_meta"

[project]
name = "ftl-reference"
version = "0.0.0"

[tool.setuptools]
py-modules = ["train", "utils", "model", "dataset"]
"""


def install_reference() -> str:
    """(Re)create baseline/_ref from /root/reference the way DESIGN.md records it: pip install --target from a copy
    under /tmp to which only a pyproject.toml was added (the reference has none).  '' on success."""
    if not all(os.path.exists(os.path.join(REF_SOURCE, f)) for f in REF_FILES):
        return f"{REF_SOURCE} is not on this machine"
    tmp = tempfile.mkdtemp(prefix="ftl_refcopy_")
    try:
        for f in REF_FILES:
            shutil.copy(os.path.join(REF_SOURCE, f), os.path.join(tmp, f))
        with open(os.path.join(tmp, "pyproject.toml"), "w") as fh:
            fh.write(_PYPROJECT)
        r = subprocess.run([sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--no-deps",
                            "--find-links", "/opt/wheelhouse", "--target", REF_DIR, tmp], capture_output=True, text=True)
        if r.returncode != 0:
            return "pip install --target baseline/_ref failed: " + (r.stderr.strip().splitlines() or ["?"])[-1][:200]
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    return ""


def reference_available() -> str:
    """'' when the reference install is usable, else the reason.  A missing install is redone when the source is here."""
    missing = [f for f in REF_FILES if not os.path.exists(os.path.join(REF_DIR, f))]
    if missing and int(os.environ.get("RANK", "0")) == 0:
        why = install_reference()
        if why:
            return f"baseline/_ref/{missing[0]} missing and {why}"
    for _ in range(150):   # other ranks wait for rank 0's install
        if all(os.path.exists(os.path.join(REF_DIR, f)) for f in REF_FILES):
            return ""
        if int(os.environ.get("RANK", "0")) == 0:
            break
        time.sleep(0.2)
    return "baseline/_ref is missing (pip install --target baseline/_ref not done)"


# ---- patching the reference's train module, re-entrantly --------------------------------------------------------------
_PRISTINE: dict = {}   # id(module) -> {"Transformer": cls, "DataLoader": cls}


def pristine_names(ref_train) -> dict:
    """The reference's own Transformer / DataLoader as they were when the module was first seen."""
    key = id(ref_train)
    if key not in _PRISTINE:
        _PRISTINE[key] = {"Transformer": ref_train.Transformer, "DataLoader": ref_train.DataLoader}
    return _PRISTINE[key]


def make_parallel_transformer(RefTransformer, mode: str, local_rank: int):
    """Subclass of the reference model that trains under torch DDP or torch FSDP2 without touching train.py."""
    import torch

    if mode == "ddp":
        class _Body(torch.nn.Module):
            def __init__(self, outer):
                super().__init__()
                self.outer = [outer]
                for n, p in outer.named_parameters():
                    self.register_parameter(n.replace(".", "__"), p)

            def forward(self, tokens):
                return RefTransformer.forward(self.outer[0], tokens)   # the pristine forward, never the subclass'

        class DDPTransformer(RefTransformer):
            def forward(self, tokens):
                from torch.nn.parallel import DistributedDataParallel as DDP

                ddp = self.__dict__.get("_ddp")
                if ddp is None:
                    ddp = DDP(_Body(self), device_ids=[local_rank], gradient_as_bucket_view=True)
                    self.__dict__["_ddp"] = ddp
                return ddp(tokens)

        return DDPTransformer

    if mode == "fsdp":
        class FSDPTransformer(RefTransformer):
            def to(self, *a, **k):   # train.py:59 `model.to(device)`: shard right there, before AdamW is built
                out = RefTransformer.to(self, *a, **k)
                if not self.__dict__.get("_fsdp_done"):
                    from torch.distributed.fsdp import fully_shard

                    for blk in self.layers.values() if hasattr(self.layers, "values") else self.layers:
                        fully_shard(blk)
                    fully_shard(self)
                    self.__dict__["_fsdp_done"] = True
                return out

        return FSDPTransformer
    raise ValueError(f"unknown parallel mode {mode!r}")


@contextlib.contextmanager
def patched(ref_train, *, transformer=None, loader=None):
    """Swap `Transformer` / `DataLoader` on the reference's train module for the duration of one call."""
    orig = pristine_names(ref_train)
    try:
        if transformer is not None:
            ref_train.Transformer = transformer
        if loader is not None:
            ref_train.DataLoader = loader
        yield orig
    finally:
        ref_train.Transformer = orig["Transformer"]
        ref_train.DataLoader = orig["DataLoader"]


_LOGGER_READY = False


def run_reference(n_gpus: int, steps: int, warmup: int, seq_len: int, batch: int, data_path: str, clock_sampler=None,
                  mode: str = "ddp", e2e: bool = True, ckpt_timing: bool = False) -> dict:
    """One call of the reference's train(args): `warmup` steps, `steps` device-timed steps at the stock logging
    frequency (5), then — if e2e — `steps` more with the loss read on the host every step (wall clock).
    With ckpt_timing (single rank or DDP only) a real SIGUSR1 is then delivered to this process inside the loop and the
    time until the reference's handle_exit has written its checkpoint is returned as `sigusr1_to_durable_s`."""
    global _LOGGER_READY
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
    orig = pristine_names(ref_train)

    transformer = None
    if world > 1:
        import torch.distributed as dist

        if not dist.is_initialized():
            dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local_rank}"))
        transformer = make_parallel_transformer(orig["Transformer"], mode, local_rank)

    def boundary():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    state = {"calls": 0, "t_e2e0": None, "t_sig": None}
    total_steps = warmup + steps * (2 if e2e else 1) + (2 if ckpt_timing else 0)
    args = argparse.Namespace(
        dataset=data_path, checkpoint_path=f"/tmp/ftl_ref_workdir/checkpoints_rank{rank}", checkpoint_id="",
        tokenizer_name_or_path="offline-shim", sequence_length=seq_len, batch_size=batch, fused_optimizer=False,
        learning_rate=5e-5, lr_warmup_steps=100, training_steps=total_steps, logging_frequency=5,
        grad_max_norm=1, model_dtype="bf16", compile=False, raise_error=False, error_step=10 ** 9)

    class TimedLoader(orig["DataLoader"]):
        def __iter__(self):
            inner = super().__iter__()

            class _It:
                def __iter__(s):
                    return s

                def __next__(s):
                    c = state["calls"]
                    if c == warmup:                               # start of the device-timed region
                        boundary()
                        if clock_sampler is not None:
                            clock_sampler.start()
                        ev[0].record()
                    elif c == warmup + steps:                     # end of it; start of the end-to-end region
                        ev[1].record()
                        boundary()
                        args.logging_frequency = 1
                        state["t_e2e0"] = time.perf_counter()
                    elif e2e and c == warmup + 2 * steps:         # only reached with ckpt_timing
                        boundary()
                        state["t_e2e1"] = time.perf_counter()
                    if ckpt_timing and c == total_steps - 1:
                        state["t_sig"] = time.perf_counter()
                        os.kill(os.getpid(), 10)                  # SIGUSR1 -> catch_SIG_exception -> handle_exit
                    state["calls"] += 1
                    return next(inner)

            return _It()

    if rank == 0 and not _LOGGER_READY and hasattr(ref_train, "init_logger"):
        ref_train.init_logger()
        _LOGGER_READY = True
    if ckpt_timing:   # the reference resubmits with os.system("sbatch ..."); there is no Slurm here, it logs the failure
        shutil.rmtree(args.checkpoint_path, ignore_errors=True)
    t_build = time.perf_counter()
    try:
        with patched(ref_train, transformer=transformer, loader=TimedLoader):
            try:
                ref_train.train(args)
            except SystemExit:
                pass
        t_exit = time.perf_counter()
        ev[2].record()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
    finally:
        if clock_sampler is not None:
            clock_sampler.stop()
        gc.collect()
        torch.cuda.empty_cache()
    if world > 1:
        torch.distributed.barrier()
    need = warmup + steps * (2 if e2e else 1)
    if state["calls"] < need:
        raise RuntimeError(f"reference loop ended early after {state['calls']} batches")
    out = {"total_s": t1 - t_build}
    out["device_ms"] = ev[0].elapsed_time(ev[1] if (e2e or ckpt_timing) else ev[2])
    if e2e:
        out["wall_ms"] = (state.get("t_e2e1", t1) - state["t_e2e0"]) * 1e3
    if ckpt_timing and state["t_sig"] is not None:
        files = os.listdir(args.checkpoint_path) if os.path.isdir(args.checkpoint_path) else []
        if files:
            out["sigusr1_to_durable_s"] = t_exit - state["t_sig"]
            out["ckpt_bytes"] = sum(os.path.getsize(os.path.join(args.checkpoint_path, f)) for f in files)
            out["ckpt_files"] = files
            # resume cost of the reference: torch.load(map_location="cpu") of that file (train.py:22) — the model /
            # optimizer load_state_dict that follow are not included
            t0 = time.perf_counter()
            ck = torch.load(os.path.join(args.checkpoint_path, files[0]), map_location="cpu")
            out["resume_load_s"] = time.perf_counter() - t0
            del ck
            shutil.rmtree(args.checkpoint_path, ignore_errors=True)
    return out
