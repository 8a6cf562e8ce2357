"""Process/rank environment and the control plane.

The reference is single-process; its only rank-related line is
`os.getenv('LOCAL_RANK', 0)` (train.py:16).  Here one process drives one GPU
(torchrun / srun style env: RANK, LOCAL_RANK, WORLD_SIZE, MASTER_ADDR/PORT).

Two process groups are created on multi-rank runs:
  * the default group — NCCL on GPUs (bootstrap, handle exchange, baseline
    collectives) or gloo on CPU;
  * `ctrl` — always gloo, CPU tensors.  Per-step scalar agreements (stop flag,
    token count) go through it so the training thread never has to synchronise
    the GPU stream just to read a flag.
"""
from __future__ import annotations

import datetime
import os
from dataclasses import dataclass
from typing import Optional

import torch
import torch.distributed as dist


@dataclass
class DistEnv:
    rank: int = 0
    local_rank: int = 0
    world_size: int = 1
    device: torch.device = torch.device("cpu")
    ctrl: Optional[object] = None  # gloo group for host-side scalars
    owns_pg: bool = False

    @property
    def is_main(self) -> bool:
        return self.rank == 0

    @property
    def distributed(self) -> bool:
        return self.world_size > 1

    # ---- control-plane scalars (host tensors over gloo) ----
    def max_int(self, value: int) -> int:
        if not self.distributed:
            return value
        t = torch.tensor([value], dtype=torch.int64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.ctrl)
        return int(t.item())

    def sum_int(self, value: int) -> int:
        if not self.distributed:
            return value
        t = torch.tensor([value], dtype=torch.int64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.ctrl)
        return int(t.item())

    def sum_float(self, value: float) -> float:
        if not self.distributed:
            return value
        t = torch.tensor([value], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.ctrl)
        return float(t.item())

    def max_float(self, value: float) -> float:
        if not self.distributed:
            return value
        t = torch.tensor([value], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.ctrl)
        return float(t.item())

    def step_sync(self, stop_code: int, n_items: int):
        """The ONE host collective of a training step: agree on a stop code and sum the token counts.

        A single SUM all-reduce of [n_items, #timeout, #error, #cancel].  Because every step performs exactly this
        collective and nothing else, a rank that fails locally can always pair with its peers' next step: it calls
        step_sync(STOP_ERROR, 0) from its exception handler (train_loop).  Returns (agreed_code, total_items);
        cancel beats error beats timeout, as in ft.signals."""
        if not self.distributed:
            return stop_code, n_items
        t = torch.zeros(4, dtype=torch.int64)
        t[0] = n_items
        if 1 <= stop_code <= 3:
            t[stop_code] = 1
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.ctrl)
        code = 3 if t[3] > 0 else (2 if t[2] > 0 else (1 if t[1] > 0 else 0))
        return code, int(t[0])

    def barrier(self) -> None:
        if self.distributed:
            dist.barrier(group=self.ctrl)

    def gather_objects(self, obj):
        if not self.distributed:
            return [obj]
        out = [None] * self.world_size
        dist.all_gather_object(out, obj, group=self.ctrl)
        return out

    def shutdown(self) -> None:
        if self.owns_pg and dist.is_initialized():
            try:
                dist.destroy_process_group()
            except Exception:
                pass


def pick_device(requested: str, local_rank: int) -> torch.device:
    if requested == "cpu":
        return torch.device("cpu")
    if requested in ("auto", "cuda"):
        if torch.cuda.is_available():
            return torch.device(f"cuda:{local_rank}")
        if requested == "cuda":
            raise RuntimeError("--device cuda requested but no CUDA device is visible")
        return torch.device("cpu")
    return torch.device(requested)


def init_dist(device_arg: str = "auto", timeout_s: int = 600) -> DistEnv:
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    device = pick_device(device_arg, local_rank)
    if device.type == "cuda":
        torch.cuda.set_device(device)
    env = DistEnv(rank=rank, local_rank=local_rank, world_size=world, device=device)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        td = datetime.timedelta(seconds=timeout_s)
        if not dist.is_initialized():
            # The data plane is this framework's own NVLink kernels (csrc/comm.cu); torch.distributed is only the
            # control plane (handle exchange, per-step scalars), so gloo suffices.  FTL_NCCL=1 brings NCCL up as
            # the default group as well (for A/B runs against library collectives).
            use_nccl = device.type == "cuda" and os.environ.get("FTL_NCCL", "0") == "1"
            kw = {"device_id": device} if use_nccl else {}
            dist.init_process_group(backend="nccl" if use_nccl else "gloo", rank=rank, world_size=world, timeout=td, **kw)
            env.owns_pg = True
        env.ctrl = dist.new_group(backend="gloo", timeout=td)
    return env
