"""Checkpoint file in the reference's layout.

One `torch.save` zip named `checkpoint_{JOBID}.ckpt` with the reference's four
keys in the reference's order (utils.py:75-80):

    model          state_dict, 291 tensors with the reference's names
    optimizer      torch AdamW state_dict (state[i] = {step, exp_avg, exp_avg_sq})
    lr_scheduler   LambdaLR state_dict
    training_step  int — number of completed steps

plus extra keys the reference's loader never touches (it indexes keys
explicitly, train.py:36-83): `data_state` (dataset position, which the
reference's README promises but never saves: SURVEY §5.4) and `meta`.

Unlike the reference the write is atomic (temp file, fsync, rename, directory
fsync) and tensors are stored as CPU tensors.
"""
from __future__ import annotations

import os
import time
from typing import Any, Dict, Optional

import torch

FORMAT_VERSION = 1


def checkpoint_file(checkpoint_path: str, jobid: str) -> str:
    return os.path.join(checkpoint_path, f"checkpoint_{jobid}.ckpt")


def _to_cpu(obj: Any) -> Any:
    if isinstance(obj, torch.Tensor):
        return obj.detach().to("cpu")
    if isinstance(obj, dict):
        return {k: _to_cpu(v) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return type(obj)(_to_cpu(v) for v in obj)
    return obj


def clean_model_state(model) -> Dict[str, torch.Tensor]:
    """state_dict with compile/DDP wrapper prefixes stripped, so resume never
    silently loads nothing the way `--compile` + strict=False does in the
    reference."""
    sd = model.state_dict()
    out = {}
    for k, v in sd.items():
        for prefix in ("_orig_mod.", "module."):
            while k.startswith(prefix):
                k = k[len(prefix):]
        out[k] = v
    return out


def atomic_torch_save(obj: Any, path: str) -> None:
    d = os.path.dirname(os.path.abspath(path))
    os.makedirs(d, exist_ok=True)
    tmp = f"{path}.tmp.{os.getpid()}"
    with open(tmp, "wb") as f:
        torch.save(obj, f)
        f.flush()
        os.fsync(f.fileno())
    os.replace(tmp, path)
    try:
        dfd = os.open(d, os.O_RDONLY)
        try:
            os.fsync(dfd)
        finally:
            os.close(dfd)
    except OSError:
        pass


def build_payload(model, optimizer, lr_scheduler, training_step: int,
                  data_state: Optional[dict] = None, extra: Optional[dict] = None) -> Dict[str, Any]:
    payload = {
        "model": _to_cpu(clean_model_state(model)),
        "optimizer": _to_cpu(optimizer.state_dict()),
        "lr_scheduler": lr_scheduler.state_dict(),
        "training_step": int(training_step),
    }
    if data_state is not None:
        payload["data_state"] = data_state
    meta = {"format_version": FORMAT_VERSION, "saved_unix_time": time.time()}
    if extra:
        meta.update(extra)
    payload["meta"] = meta
    return payload


def save_checkpoint(checkpoint_path: str, jobid: str, model, optimizer, lr_scheduler, training_step: int,
                    *, data_state: Optional[dict] = None, rank: int = 0, world_size: int = 1,
                    extra: Optional[dict] = None, logger=None) -> str:
    """Single-writer save (1 process, or replicated state with rank 0 writing)."""
    path = checkpoint_file(checkpoint_path, jobid)
    if rank == 0:
        payload = build_payload(model, optimizer, lr_scheduler, training_step, data_state, extra)
        atomic_torch_save(payload, path)
    return path


def load_checkpoint(checkpoint_path: str, checkpoint_id: str) -> Dict[str, Any]:
    """torch.load to CPU (reference: train.py:22)."""
    path = checkpoint_file(checkpoint_path, checkpoint_id)
    if not os.path.exists(path):
        # a job killed while consolidating (or while replacing an older checkpoint of the same id) leaves only
        # durable shard directories: resume from the complete one with the largest training_step
        from .sharded import best_shard_dir, load_sharded
        d = best_shard_dir(checkpoint_path, checkpoint_id)
        if d is not None:
            return load_sharded(d)
    return torch.load(path, map_location="cpu", weights_only=False)
