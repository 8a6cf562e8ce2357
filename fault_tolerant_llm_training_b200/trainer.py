"""Public training API: `make_trainer(...)` -> object with `.step(inputs, labels)`.

One step is exactly the body of the reference's loop (train.py:93-110): move the
batch to the device, zero grads, forward, sum-CE / token count, backward, grad
norm (+clip), AdamW step, LR-scheduler step.  Two implementations:

  * `TorchTrainer`  — portable PyTorch autograd path.  Runs on CPU (BASELINE.json
    config 1, gloo for multi-rank tests) and is the oracle for the native path.
  * `NativeTrainer` (models/engine.py) — B200 path: hand-scheduled forward and
    backward over the sm_100a kernels, flat bf16 parameter / gradient / moment
    buffers, fused AdamW, fused NVLink collectives on multi-GPU.

Both expose `.model`, `.optimizer`, `.lr_scheduler` with torch-compatible
state_dicts, which is what the checkpoint layout needs.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.distributed as dist

from .config import ModelConfig
from .models.llama import build_model
from .ops import torch_ops as T
from .optim.clip import clip_grad_norm_
from .optim.lr import build_lr_scheduler
from .parallel.env import DistEnv


class TorchTrainer:
    def __init__(self, cfg: ModelConfig, env: DistEnv, *, dtype: torch.dtype, lr: float, warmup_steps: int,
                 grad_max_norm: float, clip_mode: str = "clip", seed: int = 0, fused_optimizer: bool = False,
                 compile_model: bool = False):
        self.cfg, self.env = cfg, env
        self.device = env.device
        self.model = build_model(cfg, dtype, self.device, seed=seed)
        self.model.train()
        # reference: `--compile` -> torch.compile(model, fullgraph=True) (train.py:61-63).  Honoured on this portable
        # path only (the B200 engine has no tracing compiler on its hot path).  The compiled callable is kept apart
        # from `self.model`, so checkpoints keep un-prefixed names and resume stays strict.
        self._forward = torch.compile(self.model, fullgraph=True) if compile_model else self.model
        # torch.optim.AdamW with library defaults: betas (0.9, 0.999), eps 1e-8, weight_decay 0.01 on
        # every parameter, as the reference (train.py:68)
        self.optimizer = torch.optim.AdamW(self.model.parameters(), lr=lr,
                                           fused=bool(fused_optimizer and self.device.type == "cuda"))
        self.lr_scheduler = build_lr_scheduler(self.optimizer, warmup_steps)
        self.grad_max_norm, self.clip_mode = grad_max_norm, clip_mode
        self.last_grad_norm: Optional[torch.Tensor] = None
        self.kernel_launches_per_step = 0

    # reference: train.py:93-110
    def step(self, inputs: torch.Tensor, labels: torch.Tensor, n_items: Optional[int] = None) -> torch.Tensor:
        if n_items is None:  # global count of unmasked tokens; train.py passes it from its step_sync
            n_items = self.env.sum_int(int(labels.ne(-100).sum()))
        inputs = inputs.to(self.device, non_blocking=True)
        labels = labels.to(self.device, non_blocking=True)
        self.optimizer.zero_grad()
        logits = self._forward(inputs)
        loss = T.cross_entropy_sum(logits, labels) / max(n_items, 1)
        del logits
        loss.backward()
        if self.env.distributed:
            self._allreduce_grads()
            loss = loss.detach().clone()
            dist.all_reduce(loss)
        self.last_grad_norm = clip_grad_norm_(self.model.parameters(), self.grad_max_norm, self.clip_mode)
        self.optimizer.step()
        self.lr_scheduler.step()
        return loss.detach()

    def _allreduce_grads(self) -> None:
        grads = [p.grad for p in self.model.parameters() if p.grad is not None]
        flat = torch.cat([g.reshape(-1).float() for g in grads])
        dist.all_reduce(flat)
        off = 0
        for g in grads:
            n = g.numel()
            g.copy_(flat[off:off + n].view_as(g))
            off += n

    # ---- checkpoint plumbing ----
    def load_state(self, ckpt: dict) -> None:
        if "_sharded_dir" in ckpt:  # shards written by the B200 engine, resumed on the portable path
            from .ckpt.sharded import assemble_payload
            ckpt = assemble_payload(ckpt["_sharded_dir"], ckpt["_manifest"])
        self.model.load_state_dict(ckpt["model"], strict=True)  # strict
        self.optimizer.load_state_dict(ckpt["optimizer"])
        self.lr_scheduler.load_state_dict(ckpt["lr_scheduler"])

    def save_fn(self):
        return None  # default single-writer save (ckpt.layout.save_checkpoint)

    def finalize_fn(self):
        return None

    def loss_value(self, loss: torch.Tensor) -> float:
        return float(loss.item())

    def close(self, release: bool = False) -> None:
        pass


def make_trainer(cfg: ModelConfig, env: DistEnv, *, dtype: torch.dtype, lr: float, warmup_steps: int,
                 grad_max_norm: float = 1.0, clip_mode: str = "clip", seed: int = 0, parallel: str = "auto",
                 fused_optimizer: bool = False,
                 force_torch: bool = False, keep_shards: bool = False, consolidate: bool = True,
                 compile_model: bool = False):
    """Factory used by train.py, bench.py and the tests."""
    if parallel == "auto":
        parallel = "dp" if env.distributed else "none"
    if env.device.type == "cuda" and not force_torch:
        from .models.engine import NativeTrainer  # imports the sm_100a extension; fails loudly if missing

        if env.distributed and parallel == "none":
            raise ValueError("--parallel none with WORLD_SIZE > 1: ranks would train unsynchronised copies")
        if parallel == "fsdp":
            from .parallel.fsdp import FsdpTrainer as NativeTrainer  # noqa: F811
        else:
            # bucketed, overlapped reduce-scatter / AdamW / all-gather on a second stream (1..8 GPUs)
            from .parallel.pipeline import PipelinedTrainer as NativeTrainer  # noqa: F811

        return NativeTrainer(cfg, env, dtype=dtype, lr=lr, warmup_steps=warmup_steps,
                             grad_max_norm=grad_max_norm, clip_mode=clip_mode, seed=seed,
                             parallel=parallel,
                             keep_shards=keep_shards, consolidate=consolidate)
    return TorchTrainer(cfg, env, dtype=dtype, lr=lr, warmup_steps=warmup_steps, grad_max_norm=grad_max_norm,
                        clip_mode=clip_mode, seed=seed, fused_optimizer=fused_optimizer, compile_model=compile_model)
