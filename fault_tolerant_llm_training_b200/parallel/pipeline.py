"""Pipelined update engine for 1..8 GPUs: gradient reduce-scatter overlapped with backward, sharded
AdamW + parameter all-gather overlapped with the next step's forward (BASELINE.json fused path 1).

The flat parameter space is cut into buckets in forward order — embedding | layer 0 | ... | layer L-1 |
final norm + LM head — and every bucket is split into W equal sub-shards; rank r owns sub-shard r of every
bucket (so all ranks have reducible work as soon as ANY bucket's gradients exist).  A second CUDA stream
runs the data plane of csrc/comm.cu:

  backward of bucket b done (event) ──► [barrier] reduce-scatter ⊕ norm² of sub-shard (b, r)      phase 1
  all buckets reduced ──► publish {sumsq, loss} ► [barrier] ► finalize (global norm, clip, loss) ► result D2H
  for b in forward order: clip ⊕ AdamW ⊕ all-gather of sub-shard (b, r) ► [barrier] ► event "bucket b ready"
                                                                                                   phase 2
  next step's forward waits on "bucket b ready" just before the first kernel that reads bucket b.

Phase 1 hides under the backward GEMMs, phase 2 under the next forward: the HBM-bound optimizer and the
NVLink traffic run while the tensor cores are busy.  With W = 1 the same schedule (without barriers and
peer traffic) overlaps the 113 GB AdamW pass and the gradient-norm pass with compute.  Barriers are only
issued for W > 1.  Global-norm clipping needs the norm of ALL gradients before any update, which is why
the update is a second phase rather than part of the reduce-scatter kernel.
"""
from __future__ import annotations

import os
from typing import List, Tuple

import torch

from ..models.engine import FlatState, LossHandle, NativeTrainer
from ..ops import native as N
from .symm import make_symmetric


class PipelinedTrainer(NativeTrainer):
    RS_CTAS = int(os.environ.get("FTL_RS_CTAS", "96"))       # SMs' worth of CTAs lent to phase 1 while GEMMs run
    UPD_CTAS = int(os.environ.get("FTL_UPD_CTAS", "0"))      # ... and to phase 2 (0 = 4 CTAs per SM: HBM-bound)
    # The embedding bucket's gradients are the last ones of the backward: nothing else runs while its phase 1 (1 GB at
    # W = 1) delays the norm, the first AdamW launch and with them the next forward, so it gets the whole GPU.
    RS_LAST_CTAS = int(os.environ.get("FTL_RS_LAST_CTAS", "0"))
    # W > 1: the embedding bucket (the last gradient of the backward, 1 GB of mostly zeros) is reduced row-sparsely: the
    # ranks exchange their token ids and only rows that occur in some rank's batch cross NVLink (csrc/comm.cu).
    SPARSE_EMB = os.environ.get("FTL_SPARSE_EMB_RS", "1") != "0"
    TOK_CAP = 65536          # token ids per rank and step the id exchange buffer holds

    # ---- state ---------------------------------------------------------------------------------------------------
    def _make_flat_state(self) -> FlatState:
        env, W = self.env, self.env.world_size
        from ..models.llama import flat_size, param_layout
        layout = param_layout(self.model)
        numel = (flat_size(layout) + 1023) // 1024 * 1024
        flag_words = N.lib().ftl_comm_flag_words()
        # FTL_COLLECTIVE=p2p (default): fixed-order peer loads / stores, bitwise reproducible — what kill/resume chains
        # need.  FTL_COLLECTIVE=nvls: the reduction and the broadcast are done by the NVSwitch (multimem instructions on
        # a multicast mapping of p and g); fewer NVLink bytes, summation order up to the switch.
        self.nvls = os.environ.get("FTL_COLLECTIVE", "p2p") == "nvls" and W > 1
        self.symm = make_symmetric(env, dict({"p": numel * 2, "g": numel * 2, "flags": flag_words * 4, "slots": 64},
                                             **self._tok_buffers()), multicast=("p", "g") if self.nvls else ())
        bufs = {"p": lambda n: self.symm.tensor("p", torch.bfloat16), "g": lambda n: self.symm.tensor("g", torch.bfloat16)}
        flat = FlatState(self.model, self.device, buffers=bufs, pad_to=1024)
        assert flat.numel == numel
        # buckets in forward order
        L = self.cfg.n_layers
        starts = [flat.index["tok_embeddings.weight"][0]]
        starts += [flat.index[f"layers.{i}.attention.wq.weight"][0] for i in range(L)]
        starts += [flat.index["norm.weight"][0], numel]
        self.buckets: List[Tuple[int, int]] = [(starts[i], starts[i + 1]) for i in range(L + 2)]
        self.ranges = self._owned_ranges(flat)        # per bucket: flat ranges (lo, hi) this rank reduces and updates
        NB = len(self.buckets)
        self.parts = max(N.dp_partials(self.RS_CTAS), N.dp_partials(self.RS_LAST_CTAS))
        self.range_slot, k = [], 0                    # index of each range's partial-sum slice
        for rs in self.ranges:
            self.range_slot.append(list(range(k, k + len(rs))))
            k += len(rs)
        self.rs_partials = torch.zeros(max(k, 1) * self.parts, dtype=torch.float32, device=self.device)
        self.comm_status = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.commits = torch.zeros(1, dtype=torch.int64, device=self.device)
        self.slots = self.symm.tensor("slots", torch.float32)
        self.epoch = 0
        self.barrier_timeout_ms = int(os.environ.get("FTL_BARRIER_TIMEOUT_MS", "120000"))
        # high priority: when an SM frees resources, the (small, memory-bound) data-plane CTAs are placed first
        self.comm = torch.cuda.Stream(device=self.device, priority=-1)
        self.ev_grad = [torch.cuda.Event() for _ in range(NB)]
        self.ev_param = [torch.cuda.Event() for _ in range(NB)]
        self.ev_step = torch.cuda.Event()
        self._params_pending = False
        self._hyper = None
        return flat

    def _tok_buffers(self) -> dict:
        """Peer-mapped buffer for the token-id exchange of the sparse embedding reduce: 3 step slots x W ranks x TOK_CAP."""
        W = self.env.world_size
        return {"toks": 3 * W * self.TOK_CAP * 8} if (W > 1 and self.SPARSE_EMB) else {}

    def _owned_ranges(self, flat: FlatState) -> List[List[Tuple[int, int]]]:
        """Data parallel: every bucket is cut into W equal contiguous sub-shards; rank r owns the r-th."""
        W, r, out = self.env.world_size, self.env.rank, []
        for lo, hi in self.buckets:
            sz = ((hi - lo + W - 1) // W + 7) // 8 * 8
            a = min(lo + r * sz, hi)
            e = min(a + sz, hi)
            out.append([(a, e)] if e > a else [])
        return out

    def _moment_ranges(self):
        return [rg for rs in self.ranges for rg in rs]

    # ---- hooks called from the engine ----------------------------------------------------------------------------
    def _ensure_engine(self, B: int, S: int) -> None:
        fresh = self.engine is None or (self.engine.B, self.engine.S) != (B, S)
        super()._ensure_engine(B, S)
        if fresh:
            self.engine.hook_params = self._wait_params
            self.engine.hook_grads = self._grads_done
            W, T = self.env.world_size, B * S
            can = (W > 1 and self.SPARSE_EMB and T <= self.TOK_CAP and len(self.ranges[0]) == 1
                   and "toks" in self.symm.sizes)
            # every rank must take the same path (a rank that owns no part of a very small table would not qualify)
            self.sparse_emb = bool(W > 1 and self.env.max_int(0 if can else 1) == 0)
            if self.sparse_emb:
                self.engine.hook_tokens = self._push_tokens
                self.engine.sparse_emb_zero = True       # the backward clears only its own previous rows: no 1 GB memset
                self.engine.g_emb.zero_()                # once: from here on every row outside the token lists stays zero
                self.toks_local = self.symm.tensor("toks", torch.int64)
                self._tok_step, self._have_prev = 0, False
                k = sum(len(rs) for rs in self.ranges)
                self.rs_partials = torch.zeros(max(k, 1) * self.parts + W * T, dtype=torch.float32, device=self.device)
                torch.cuda.synchronize()
                self.env.barrier()
            # One rank has nothing to reduce: the kernels that write the gradients also emit the norm partials, and
            # phase 1 (a 16 GB re-read of the gradient buffer at W = 1) disappears.
            self.fused_norm = (type(self) is PipelinedTrainer and self.env.world_size == 1
                               and os.environ.get("FTL_FUSED_NORM", "1") != "0" and self.engine.enable_fused_norm())

    def _wait_params(self, b: int) -> None:
        """Forward is about to read bucket b: wait until last step's update of it has landed everywhere."""
        if self._params_pending:
            torch.cuda.current_stream().wait_event(self.ev_param[b])

    def _barrier(self) -> None:
        if self.env.world_size > 1:
            self.epoch += 1
            N.barrier(self.symm.table("flags"), self.env.rank, self.env.world_size, self.epoch, self.comm_status,
                      self.barrier_timeout_ms)

    def _grads_done(self, b: int) -> None:
        """Backward finished writing bucket b's gradients: reduce this rank's sub-shard on the comm stream."""
        self.ev_grad[b].record()
        with torch.cuda.stream(self.comm):
            self.comm.wait_event(self.ev_grad[b])
            if getattr(self, "fused_norm", False):
                return
            self._barrier()          # every rank's bucket-b gradients are complete
            if b == 0 and getattr(self, "sparse_emb", False):
                self._reduce_embedding_sparse()
                return
            for (a, e), k in zip(self.ranges[b], self.range_slot[b]):
                part = self.rs_partials[k * self.parts:(k + 1) * self.parts]
                ctas = self.RS_LAST_CTAS if b == 0 else self.RS_CTAS
                if getattr(self, "nvls", False):
                    N.nvls_reduce_scatter(self.symm.mc_ptr["g"], self.flat.g, a, e - a, part, self.comm_status, max_ctas=ctas)
                else:
                    N.dp_reduce_scatter(self.symm.table("g"), self.env.rank, self.env.world_size, a, e - a, part,
                                        self.comm_status, max_ctas=ctas)

    def _push_tokens(self, flat_tok: torch.Tensor) -> None:
        """Start of the backward (compute stream): hand this step's token ids to every peer."""
        W, cap = self.env.world_size, self.TOK_CAP
        self._tok_slot = self._tok_step % 3
        N.emb_tokens_push(flat_tok, self.symm.table("toks"), self._tok_slot * W * cap, self.env.rank, W, cap)

    def _reduce_embedding_sparse(self) -> None:
        """Comm stream, after the barrier that follows the embedding backward on every rank."""
        W, cap, e = self.env.world_size, self.TOK_CAP, self.engine
        cur = self.toks_local[self._tok_slot * W * cap:]
        prev = self.toks_local[((self._tok_step - 1) % 3) * W * cap:] if self._have_prev else None
        table, emb_off, a, hi = self._emb_grad_location()
        tail = self.rs_partials[self.rs_partials.numel() - W * e.T:]
        N.emb_sparse_reduce(table, self.env.rank, W, cur, prev, e.T, cap, emb_off, self.cfg.dim,
                            self.cfg.vocab_size, a, hi, tail, self.comm_status)
        self._tok_step += 1
        self._have_prev = True

    def _emb_grad_location(self):
        """(peer table of the buffer holding the embedding gradient, element offset of the table in it, this rank's
        element range [lo, hi) in it)."""
        (a, hi), = self.ranges[0]
        return self.symm.table("g"), self.flat.index["tok_embeddings.weight"][0], a, hi

    # ---- step pieces ---------------------------------------------------------------------------------------------
    def _begin_step(self, hyper: torch.Tensor) -> None:
        self._hyper = hyper
        # the comm stream must not start this step's phase 1 before the staging copy (hyper) has been issued
        self.ev_step.record()
        self.comm.wait_event(self.ev_step)

    def _reduce_and_update(self, hyper: torch.Tensor, loss: torch.Tensor) -> None:
        env, f, W = self.env, self.flat, self.env.world_size
        with torch.cuda.stream(self.comm):
            parts = self.engine.norm_arena if getattr(self, "fused_norm", False) else self.rs_partials
            N.dp_publish(parts, parts.numel(), loss, self.symm.table("slots"), env.rank, W)
            self._barrier()          # all slots written; nobody still reads my gradients for phase 1
            N.dp_finalize(self.slots, W, hyper, self.stats, loss, clip=(self.clip_mode == "clip"), status=self.comm_status,
                          result=self.result_dev)
            self._result_handle = self._publish_result_on_comm(loss)
            for b in range(len(self.buckets)):
                self._update_bucket(b, hyper)
                self._barrier()      # bucket b is updated on every rank
                self.ev_param[b].record()
                self._after_update(b)
        self._params_pending = True
        self.optimizer.steps_done += 1
        self.optimizer._opt_called = True   # the update ran inside the fused kernels

    def _update_bucket(self, b: int, hyper: torch.Tensor) -> None:
        """clip ⊕ AdamW ⊕ all-gather of this rank's ranges of bucket b (runs on the comm stream)."""
        f, env = self.flat, self.env
        for a, e in self.ranges[b]:
            if getattr(self, "nvls", False):
                N.nvls_adamw_allgather(self.symm.mc_ptr["p"], f.p, f.g, f.m, f.v, a, e - a, hyper, self.stats,
                                       self.comm_status, self.commits, max_ctas=self.UPD_CTAS)
            else:
                N.dp_adamw_allgather(self.symm.table("p"), f.g, f.m, f.v, env.rank, env.world_size, a, e - a, hyper,
                                     self.stats, self.comm_status, self.commits, max_ctas=self.UPD_CTAS)

    def _after_update(self, b: int) -> None:
        pass

    def _publish_result_on_comm(self, loss: torch.Tensor) -> LossHandle:
        slot = self._result_slot       # result_dev was filled by the finalize kernel: one D2H copy, no packing kernels
        self.result_host[slot].copy_(self.result_dev, non_blocking=True)
        self.events[slot].record()
        return LossHandle(self.result_host[slot], self.events[slot])

    def _publish_result(self, loss: torch.Tensor, slot: int) -> LossHandle:
        return self._result_handle      # recorded on the comm stream right after finalize

    def step(self, inputs, labels, n_items=None):
        self._result_slot = self._slot   # ring slot this step will use (base class advances _slot)
        return super().step(inputs, labels, n_items=n_items)

    # ---- fault tolerance -----------------------------------------------------------------------------------------
    def committed_updates(self) -> int:
        """Number of optimizer steps fully applied on the device (each applies one update per non-empty bucket)."""
        torch.cuda.synchronize()
        per_step = sum(len(rs) for rs in self.ranges)
        return int(self.commits.item()) // max(per_step, 1)

    def abort_peers(self) -> None:
        """A rank that leaves the loop on its own first lets its already-enqueued steps finish (the peers still take
        part in those), then raises the abort word so that peers' kernels waiting for a step this rank will never
        run fall through their barriers with status 2 instead of spinning until the timeout."""
        if self.env.world_size > 1:
            try:
                torch.cuda.synchronize()
            except RuntimeError:
                pass
            N.raise_abort(self.symm.table("flags"), self.env.world_size)
            torch.cuda.synchronize()

    def load_state(self, ckpt: dict) -> None:
        super().load_state(ckpt)
        torch.cuda.synchronize()
        self.env.barrier()

    def close(self, release: bool = False) -> None:
        """release=True also frees the peer-mapped buffers (a collective; for processes that build another trainer
        afterwards, e.g. bench.py).  The training driver leaves that to process exit: after a failure a collective
        in the exit path could wait for a rank that is gone."""
        torch.cuda.synchronize()
        if release:
            self.symm.close()
