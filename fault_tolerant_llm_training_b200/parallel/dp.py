"""Data-parallel engine (ZeRO-1 style) over the in-kernel NVLink data plane of csrc/comm.cu.

Every rank runs the full forward/backward on its own batch (per-rank batch fixed: weak scaling).  Then, on
the compute stream, with no host synchronisation and no NCCL call:

    barrier                       every rank's gradients are complete
    reduce-scatter ⊕ norm²        rank r sums shard r of all W gradient buffers through peer loads
    publish {sumsq, loss} to all  16 bytes per peer
    barrier                       slots visible everywhere; nobody still reads my gradients
    finalize                      identical global norm / clip coefficient / loss on every rank
    clip ⊕ AdamW ⊕ all-gather     rank r updates shard r and stores the new bf16 parameters into every
                                  rank's parameter buffer through peer stores
    barrier                       parameters consistent; the next forward may start

Optimizer moments are only maintained for the local shard (the full-size buffers exist so that checkpoints
keep torch's per-parameter layout; `sync_optimizer_state()` all-gathers them before a save).  The reference
has no multi-GPU path; the baseline this is measured against is torch DDP + NCCL
(baseline/ref_harness.py).
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from ..models.engine import FlatState, NativeTrainer
from ..ops import native as N
from .symm import SymmetricBuffers


class DPNativeTrainer(NativeTrainer):
    def _make_flat_state(self) -> FlatState:
        env = self.env
        W = env.world_size
        from ..models.llama import flat_size, param_layout
        pad = W * 1024
        numel = (flat_size(param_layout(self.model)) + pad - 1) // pad * pad
        self.shard = numel // W
        self.lo = env.rank * self.shard
        flag_words = N.lib().ftl_comm_flag_words()
        self.symm = SymmetricBuffers(env, {"p": numel * 2, "g": numel * 2, "flags": flag_words * 4, "slots": 64})
        self.comm_status = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.rs_partials = torch.zeros(N.dp_partials(), dtype=torch.float32, device=self.device)
        self.epoch = 0
        self.commits = torch.zeros(1, dtype=torch.int64, device=self.device)
        self.barrier_timeout_ms = 120000
        bufs = {"p": lambda n: self.symm.tensor("p", torch.bfloat16), "g": lambda n: self.symm.tensor("g", torch.bfloat16)}
        flat = FlatState(self.model, self.device, buffers=bufs, pad_to=pad)
        assert flat.numel == numel
        self.slots = self.symm.tensor("slots", torch.float32)
        return flat

    def _barrier(self) -> None:
        self.epoch += 1
        N.barrier(self.symm.table("flags"), self.env.rank, self.env.world_size, self.epoch, self.comm_status,
                  self.barrier_timeout_ms)

    def _reduce_and_update(self, hyper: torch.Tensor, loss: torch.Tensor) -> None:
        env, f, W = self.env, self.flat, self.env.world_size
        self._barrier()
        N.dp_reduce_scatter(self.symm.table("g"), env.rank, W, self.lo, self.shard, self.rs_partials, self.comm_status)
        N.dp_publish(self.rs_partials, self.rs_partials.numel(), loss, self.symm.table("slots"), env.rank, W)
        self._barrier()
        N.dp_finalize(self.slots, W, hyper, self.stats, loss, clip=(self.clip_mode == "clip"))
        N.dp_adamw_allgather(self.symm.table("p"), f.g, f.m, f.v, env.rank, W, self.lo, self.shard, hyper, self.stats,
                             self.comm_status, self.commits)
        self._barrier()
        self.optimizer.steps_done += 1
        self.optimizer._opt_called = True   # the update ran inside the fused kernel, not through optimizer.step()

    def _write_status(self, dst: torch.Tensor) -> None:
        dst.copy_(self.comm_status.to(torch.float32))

    def abort_peers(self) -> None:
        """Called by a rank that failed mid-step: peers spinning in a barrier fall through with status 2."""
        N.raise_abort(self.symm.table("flags"), self.env.world_size)
        torch.cuda.synchronize()

    def sync_optimizer_state(self) -> None:
        """All-gather the moment shards so that every rank (rank 0 in particular) holds full exp_avg / exp_avg_sq."""
        torch.cuda.synchronize()
        for buf in (self.flat.m, self.flat.v):
            dist.all_gather_into_tensor(buf, buf[self.lo:self.lo + self.shard].clone())
        torch.cuda.synchronize()

    def _moment_ranges(self):
        return [(self.lo, self.lo + self.shard)]

    def load_state(self, ckpt: dict) -> None:
        super().load_state(ckpt)
        torch.cuda.synchronize()
        self.env.barrier()
