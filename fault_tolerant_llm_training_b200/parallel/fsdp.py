"""Fully-sharded engine (ZeRO-3 style): BASELINE.json fused path 2 — the parameter all-gather is pulled into
the QKV and gate/up GEMMs that consume the gathered weights.

Persistent state per rank is only ITS row shard of every weight matrix (one compact peer-mapped buffer), the
matching AdamW moments and the reduced gradient shards; gradients are produced into two layer-sized slots that the
backward reuses (plus one buffer each for the embedding and the head bucket), never into a model-sized buffer.  Gathered weights live in staging memory, bucket by
bucket (bucket = embedding | layer i | final norm + head), in one of two modes:

  resident (default)   one staging buffer the size of the model (16 GB of the 180 GB HBM for the 8 B model): a bucket
                       gathered for the forward is still there when the backward needs it, so every parameter
                       crosses NVLink ONCE per step — what torch FSDP calls reshard_after_forward=False
  reshard (FTL_FSDP_RESHARD=1)   two bucket-sized slots; the backward re-gathers each bucket one bucket ahead of
                       the GEMMs that read it (twice the NVLink bytes, 12 GB less resident memory)

  forward, layer i   wo, w2 and the norm weights of the layer are gathered by the plain all-gather kernel on the
                     comm stream while the previous layer computes;  wq|wk|wv and w1, w3 — 80 % of the layer's
                     bytes — are NOT gathered ahead: `ftl_gemm_qkv_rope_gather` / `ftl_gemm_swiglu_gather`
                     pull the owners' rows over NVLink from inside the GEMM (the epilogue warps of every CTA
                     move 16-row slices while they wait for accumulators; the TMA producers wait per 128-row
                     half block), so the transfer overlaps the MMAs tile by tile
  backward, layer i  resident mode: nothing to fetch.  reshard mode: the whole bucket is re-gathered (plain kernel,
                     one bucket ahead of the backward GEMMs)
  update             gradient reduce-scatter ⊕ norm² per shard under backward, then clip ⊕ AdamW on the compact
                     shards; no parameter all-gather follows — the next forward reads the shards directly

The schedule, barriers, abort/commit logic and the sharded checkpoint writer are those of the pipelined
data-parallel engine (parallel/pipeline.py); only the ownership map and the parameter path differ.
The reference has no counterpart.
"""
from __future__ import annotations

import os
from typing import Dict, List, Tuple

import torch

from ..models.engine import FlatState
from ..ops import native as N
from .pipeline import PipelinedTrainer
from .symm import make_symmetric


class _Unit:
    """One row-sharded matrix: flat range [off, off + numel), rows x cols, `shard` elements per rank."""
    __slots__ = ("name", "off", "numel", "rows", "cols", "shard", "coff", "bucket", "fused")

    def __init__(self, name, off, numel, rows, cols, shard, coff, bucket, fused):
        self.name, self.off, self.numel, self.rows, self.cols = name, off, numel, rows, cols
        self.shard, self.coff, self.bucket, self.fused = shard, coff, bucket, fused


class FsdpState(FlatState):
    """Parameters are views into the staging memory: `slot_of(bucket)` returns the staging tensor that holds the bucket
    and the flat offset of its first element.  Gradients are views into the flat gradient buffer, or — with sharded
    gradients — into the buffer `gslot_of(bucket)` names the same way."""

    def bind_slots(self, slot_of, bucket_of: Dict[str, int], gslot_of=None) -> None:
        self._slot_of, self._bucket_of, self._gslot_of = slot_of, bucket_of, gslot_of

    def grad_view(self, name: str) -> torch.Tensor:
        if self._gslot_of is None:
            return super().grad_view(name)
        off, n, shape = self.index[name]
        buf, base = self._gslot_of(self._bucket_of[name])
        return buf[off - base:off - base + n].view(shape)

    def grad_span(self, first: str, last: str, shape) -> torch.Tensor:
        if self._gslot_of is None:
            return super().grad_span(first, last, shape)
        o0, _, _ = self.index[first]
        o1, n1, _ = self.index[last]
        buf, base = self._gslot_of(self._bucket_of[first])
        return buf[o0 - base:o1 + n1 - base].view(shape)

    def param_view(self, name: str) -> torch.Tensor:
        off, n, shape = self.index[name]
        buf, base = self._slot_of(self._bucket_of[name])
        return buf[off - base:off - base + n].view(shape)

    def param_span(self, first: str, last: str, shape) -> torch.Tensor:
        o0, _, _ = self.index[first]
        o1, n1, _ = self.index[last]
        buf, base = self._slot_of(self._bucket_of[first])
        return buf[o0 - base:o1 + n1 - base].view(shape)


class FsdpTrainer(PipelinedTrainer):
    GATHER_CTAS = 64

    # ---- state ---------------------------------------------------------------------------------------------------
    def _make_flat_state(self) -> FlatState:
        env, W, L = self.env, self.env.world_size, self.cfg.n_layers
        from ..models.llama import flat_size, param_layout
        layout = param_layout(self.model)
        numel = (flat_size(layout) + 1023) // 1024 * 1024
        index = {name: (off, n, shape) for name, shape, off, n in layout}
        # ---- buckets (forward order) and shard units
        starts = [index["tok_embeddings.weight"][0]] + [index[f"layers.{i}.attention.wq.weight"][0] for i in range(L)]
        starts += [index["norm.weight"][0], numel]
        self.buckets = [(starts[i], starts[i + 1]) for i in range(L + 2)]
        d = self.cfg.dim
        self.units: List[_Unit] = []
        self.bucket_units: List[List[_Unit]] = [[] for _ in range(L + 2)]
        coff = 0

        def add(name, off, rows, cols, bucket, fused=False):
            nonlocal coff
            n = rows * cols
            if n % (W * 8):
                raise ValueError(f"{name}: {n} elements do not split into {W} 16-byte-aligned shards")
            u = _Unit(name, off, n, rows, cols, n // W, coff, bucket, fused)
            coff += (u.shard + 127) // 128 * 128
            self.units.append(u)
            self.bucket_units[bucket].append(u)
            return u

        def mat(name, bucket, fused=False):
            off, n, shape = index[name]
            return add(name, off, shape[0] if len(shape) == 2 else 1, shape[1] if len(shape) == 2 else shape[0], bucket, fused)

        mat("tok_embeddings.weight", 0)
        self.layer_units = []
        for i in range(L):
            b, pre = i + 1, f"layers.{i}."
            oq = index[pre + "attention.wq.weight"][0]
            ov, nv, _ = index[pre + "attention.wv.weight"]
            qkv_rows = (ov + nv - oq) // d
            fusable = (qkv_rows // W) % 128 == 0 and (self.cfg.ffn_hidden // W) % 128 == 0 and W > 0
            lu = {"qkv": add(pre + "attention.wqkv", oq, qkv_rows, d, b, fused=fusable),
                  "wo": mat(pre + "attention.wo.weight", b),
                  "w1": mat(pre + "feed_forward.w1.weight", b, fused=fusable),
                  "w2": mat(pre + "feed_forward.w2.weight", b),
                  "w3": mat(pre + "feed_forward.w3.weight", b, fused=fusable),
                  "an": mat(pre + "attention_norm.weight", b), "fn": mat(pre + "ffn_norm.weight", b)}
            self.layer_units.append(lu)
        mat("norm.weight", L + 1)
        mat("output.weight", L + 1)
        self.own_numel = (coff + 1023) // 1024 * 1024
        max_bucket = max(hi - lo for lo, hi in self.buckets)
        flag_words = N.lib().ftl_comm_flag_words()
        # fully-sharded mode has no parameter broadcast; its gradient reduce-scatter can go through the switch too
        self.nvls = os.environ.get("FTL_COLLECTIVE", "p2p") == "nvls" and W > 1
        # Sharded gradients (default; FTL_FSDP_SHARD_GRADS=0 keeps a model-sized buffer): the embedding and the head bucket keep their own
        # gradient buffers, the layers write into two layer-sized slots by bucket parity, and every reduce-scatter leaves
        # its result in a compact shard buffer; a slot is reused by the backward two layers later, once every rank has
        # finished reading it (a barrier after the reduce-scatter + an event the compute stream waits on).
        # (the NVLS reduce-scatter needs one multicast mapping of the whole gradient buffer: it keeps the full buffer)
        self.shard_grads = os.environ.get("FTL_FSDP_SHARD_GRADS", "1") != "0" and W > 1 and not self.nvls
        if self.shard_grads:
            layer_max = max(hi - lo for lo, hi in self.buckets[1:L + 1])
            gsizes = {"g0": (self.buckets[0][1] - self.buckets[0][0]) * 2,
                      "gL": (self.buckets[L + 1][1] - self.buckets[L + 1][0]) * 2, "gs0": layer_max * 2, "gs1": layer_max * 2}
        else:
            gsizes = {"g": numel * 2}
        self.symm = make_symmetric(env, dict({"own": self.own_numel * 2, "flags": flag_words * 4, "slots": 64}, **gsizes,
                                             **self._tok_buffers()), multicast=("g",) if self.nvls else ())
        self.p_own = self.symm.tensor("own", torch.bfloat16)
        self.m_own = torch.zeros(self.own_numel, dtype=torch.bfloat16, device=self.device)
        self.v_own = torch.zeros(self.own_numel, dtype=torch.bfloat16, device=self.device)
        self.reshard = os.environ.get("FTL_FSDP_RESHARD", "0") == "1"
        if self.reshard:   # two bucket-sized slots, by bucket parity
            self.stage = [torch.zeros(max_bucket, dtype=torch.bfloat16, device=self.device) for _ in range(2)]
        else:              # the whole model stays gathered from its forward until its update
            self.stage = [torch.zeros(numel, dtype=torch.bfloat16, device=self.device)]
        # ---- move the freshly initialised parameters into the shards, then drop the full copies
        with torch.no_grad():
            params = dict(self.model.named_parameters())
            full = {}
            for u in self.units:
                if u.name.endswith("attention.wqkv"):
                    pre = u.name[:-len("wqkv")]
                    src = torch.cat([params[pre + k + ".weight"].reshape(-1) for k in ("wq", "wk", "wv")])
                else:
                    src = params[u.name].reshape(-1)
                self.p_own[u.coff:u.coff + u.shard].copy_(src[env.rank * u.shard:(env.rank + 1) * u.shard])
            del full
        if self.shard_grads:
            self.gbufs = {k: self.symm.tensor(k, torch.bfloat16) for k in ("g0", "gL", "gs0", "gs1")}
            self.g_own = torch.zeros(self.own_numel, dtype=torch.bfloat16, device=self.device)
            self.ev_gfree = [torch.cuda.Event(), torch.cuda.Event()]
        bufs = {"p": lambda n: torch.zeros(0, dtype=torch.bfloat16, device=self.device),   # no full parameter buffer
                "g": lambda n: (torch.zeros(0, dtype=torch.bfloat16, device=self.device) if self.shard_grads
                                else self.symm.tensor("g", torch.bfloat16)),
                "m": lambda n: torch.zeros(0, dtype=torch.bfloat16, device=self.device),
                "v": lambda n: torch.zeros(0, dtype=torch.bfloat16, device=self.device)}
        flat = FsdpState.__new__(FsdpState)
        flat.layout, flat.numel = layout, numel
        flat.p, flat.g, flat.m, flat.v = bufs["p"](0), bufs["g"](numel), bufs["m"](0), bufs["v"](0)
        flat.index = index
        bucket_of = {}
        for name, (off, n, shape) in index.items():
            bucket_of[name] = max(b for b, (lo, hi) in enumerate(self.buckets) if lo <= off)
        flat.bind_slots(self._slot_of, bucket_of, self._gslot_of if self.shard_grads else None)
        with torch.no_grad():
            for name, prm in self.model.named_parameters():
                off, n, shape = index[name]
                prm.data = torch.empty(0, dtype=torch.bfloat16, device=self.device)   # full weights are never resident
                prm.grad = None
        self._finish_pipeline_state(flat)
        # ---- staging-slot bookkeeping
        NB = L + 2
        ns = 2 if self.reshard else NB
        self.resident = [-1] * ns            # bucket currently in each slot
        self.resident_full = [False] * ns    # ... including the matrices the fused GEMMs gather themselves
        self.ev_gather = [torch.cuda.Event() for _ in range(NB)]
        self.ev_slot_free = [torch.cuda.Event() for _ in range(ns)]
        self.gather_stream = torch.cuda.Stream(device=self.device, priority=-1)
        self.gather_scratch = torch.zeros(1 + 2 * (self.cfg.ffn_hidden // 128 + self.cfg.qkv_dim // 128 + 2),
                                          dtype=torch.int32, device=self.device)
        return flat

    def _finish_pipeline_state(self, flat: FlatState) -> None:
        """The part of PipelinedTrainer._make_flat_state that does not allocate parameter buffers."""
        self.ranges = self._owned_ranges(flat)
        NB = len(self.buckets)
        self.parts = N.dp_partials(self.RS_CTAS)
        self.range_slot, k = [], 0
        for rs in self.ranges:
            self.range_slot.append(list(range(k, k + len(rs))))
            k += len(rs)
        self.rs_partials = torch.zeros(max(k, 1) * self.parts, dtype=torch.float32, device=self.device)
        self.comm_status = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.commits = torch.zeros(1, dtype=torch.int64, device=self.device)
        self.slots = self.symm.tensor("slots", torch.float32)
        self.epoch = 0
        self.barrier_timeout_ms = int(os.environ.get("FTL_BARRIER_TIMEOUT_MS", "120000"))
        self.comm = torch.cuda.Stream(device=self.device, priority=-1)
        self.ev_grad = [torch.cuda.Event() for _ in range(NB)]
        self.ev_param = [torch.cuda.Event() for _ in range(NB)]
        self.ev_step = torch.cuda.Event()
        self._params_pending = False
        self._hyper = None

    def _k(self, b: int) -> int:
        """Staging slot index of bucket b."""
        return b % 2 if self.reshard else b

    def _slot_of(self, b: int):
        """(staging tensor, flat offset of its first element) for bucket b."""
        return (self.stage[b % 2], self.buckets[b][0]) if self.reshard else (self.stage[0], 0)

    def _gname(self, b: int) -> str:
        """Name of the gradient buffer that bucket b's gradients are written to (sharded gradients)."""
        L = self.cfg.n_layers
        return "g0" if b == 0 else ("gL" if b == L + 1 else f"gs{b % 2}")

    def _gslot_of(self, b: int):
        return self.gbufs[self._gname(b)], self.buckets[b][0]

    def _emb_grad_location(self):
        if not self.shard_grads:
            return super()._emb_grad_location()
        u, r = self.bucket_units[0][0], self.env.rank
        base = self.buckets[0][0]
        return self.symm.table("g0"), u.off - base, u.off - base + r * u.shard, u.off - base + (r + 1) * u.shard

    def _grads_done(self, b: int) -> None:
        if not self.shard_grads:
            return super()._grads_done(b)
        env, L = self.env, self.cfg.n_layers
        self.ev_grad[b].record()
        with torch.cuda.stream(self.comm):
            self.comm.wait_event(self.ev_grad[b])
            self._barrier()          # every rank's bucket-b gradients are complete
            if b == 0 and getattr(self, "sparse_emb", False):
                self._reduce_embedding_sparse()      # in place in the embedding's own gradient buffer
                return
            table, base = self.symm.table(self._gname(b)), self.buckets[b][0]
            for u, k in zip(self.bucket_units[b], self.range_slot[b]):
                part = self.rs_partials[k * self.parts:(k + 1) * self.parts]
                lo = u.off - base + env.rank * u.shard
                ctas = self.RS_LAST_CTAS if b == 0 else self.RS_CTAS
                if b == 0:           # dense fallback for the embedding: in place, as in the unsharded layout
                    N.dp_reduce_scatter(table, env.rank, env.world_size, lo, u.shard, part, self.comm_status, max_ctas=ctas)
                else:
                    N.dp_reduce_scatter_to(table, env.rank, env.world_size, lo, u.shard, self.g_own[u.coff:u.coff + u.shard],
                                           part, self.comm_status, max_ctas=ctas)
            if 1 <= b <= L:
                self._barrier()      # every rank has finished reading this slot: the backward may write it again
                self.ev_gfree[b % 2].record()

    def _owned_ranges(self, flat) -> List[List[Tuple[int, int]]]:
        r = self.env.rank
        return [[(u.off + r * u.shard, u.off + (r + 1) * u.shard) for u in us] for us in self.bucket_units]

    # ---- parameter path: gathers --------------------------------------------------------------------------------
    def _gather(self, b: int, include_fused: bool) -> None:
        """Enqueue (comm stream) the plain all-gather of bucket b into staging slot b % 2, unless it is there already.
        The caller guarantees that every kernel reading the slot's previous occupant has been enqueued on the main
        stream: an event recorded there now is what the gather waits for."""
        k = self._k(b)
        have = self.resident[k] == b
        if have and (self.resident_full[k] or not include_fused):
            return
        env = self.env
        buf, lo = self._slot_of(b)
        if not have:
            self.ev_slot_free[k].record()                      # main stream: previous occupant no longer needed
        # Gathers run on their own stream: on the comm stream they would queue behind the AdamW kernels of EVERY
        # bucket of the previous step, and the forward could not start before the whole update had finished.
        with torch.cuda.stream(self.gather_stream):
            if not have:
                self.gather_stream.wait_event(self.ev_slot_free[k])
            if self._params_pending:
                self.gather_stream.wait_event(self.ev_param[b])   # bucket b's shards are updated on every rank
            for u in self.bucket_units[b]:
                if (u.fused and not include_fused) or (have and not u.fused):
                    continue
                N.fsdp_gather(self.symm.table("own"), u.coff, buf[u.off - lo:u.off - lo + u.numel], u.shard,
                              env.rank, env.world_size, self.comm_status, max_ctas=self.GATHER_CTAS)
            self.ev_gather[b].record()
        self.resident[k] = b
        self.resident_full[k] = include_fused or not any(u.fused for u in self.bucket_units[b])

    def _wait_params(self, b: int) -> None:
        """Forward: the engine is about to read bucket b (all kernels of buckets < b are already enqueued)."""
        main = torch.cuda.current_stream()
        fused = self.engine is not None and self.engine.fused_gather is not None
        self._gather(b, include_fused=not fused)      # no-op when the prefetch below already issued it
        main.wait_event(self.ev_gather[b])
        if self._params_pending:
            main.wait_event(self.ev_param[b])         # the fused GEMMs read bucket b's shards on the peers directly
        if b + 1 < len(self.buckets):
            self._gather(b + 1, include_fused=not fused)   # prefetch into the other slot while bucket b computes

    def _params_for_backward(self, b: int) -> None:
        """Backward: the engine is about to read bucket b (all kernels of buckets > b are already enqueued)."""
        main = torch.cuda.current_stream()
        if self.shard_grads and 1 <= b <= self.cfg.n_layers - 2:
            main.wait_event(self.ev_gfree[b % 2])     # bucket b + 2's gradients have been reduced out of this slot everywhere
        self._gather(b, include_fused=True)
        main.wait_event(self.ev_gather[b])
        if b - 1 >= 1:
            self._gather(b - 1, include_fused=True)   # prefetch the next (lower) layer

    def _ensure_engine(self, B: int, S: int) -> None:
        fresh = self.engine is None or (self.engine.B, self.engine.S) != (B, S)
        super()._ensure_engine(B, S)
        if fresh:
            self.engine.hook_params_bwd = self._params_for_backward
            if all(lu["qkv"].fused for lu in self.layer_units) and self.engine.USE_TC and N.has("ftl_gemm_qkv_rope_gather"):
                self.engine.fused_gather = self

    # gather-fused GEMMs, called by the engine in the forward of layer i (main stream)
    def qkv(self, i: int, x: torch.Tensor, out: torch.Tensor) -> None:
        c, u, e = self.cfg, self.layer_units[i]["qkv"], self.engine
        N.gemm_qkv_rope_gather(x, e.lw[i]["wqkv"], out, e.cos, e.sin, e.S, c.head_dim, (c.n_heads + c.n_kv_heads) * c.head_dim,
                               self.symm.table("own"), u.coff, u.rows // self.env.world_size, self.env.rank,
                               self.env.world_size, self.gather_scratch)

    def swiglu(self, i: int, x: torch.Tensor, act: torch.Tensor, gate: torch.Tensor, up: torch.Tensor) -> None:
        lu, e = self.layer_units[i], self.engine
        N.gemm_swiglu_gather(x, e.lw[i]["ww1"], e.lw[i]["ww3"], act, gate, up, self.symm.table("own"), lu["w1"].coff,
                             lu["w3"].coff, lu["w1"].rows // self.env.world_size, self.env.rank, self.env.world_size,
                             self.gather_scratch)
        self.resident_full[self._k(i + 1)] = True     # wq|wk|wv, w1, w3 of this layer are now in the slot too

    # ---- update: AdamW on the compact shards, no all-gather -----------------------------------------------------
    def _update_bucket(self, b: int, hyper: torch.Tensor) -> None:
        r = self.env.rank
        for u in self.bucket_units[b]:
            lo = u.off + r * u.shard
            if not self.shard_grads:
                g = self.flat.g[lo:lo + u.shard]
            elif b == 0:             # reduced in place in the embedding's own buffer
                g = self.gbufs["g0"][lo - self.buckets[0][0]:lo - self.buckets[0][0] + u.shard]
            else:
                g = self.g_own[u.coff:u.coff + u.shard]
            N.adamw_shard(self.p_own[u.coff:u.coff + u.shard], g,
                          self.m_own[u.coff:u.coff + u.shard], self.v_own[u.coff:u.coff + u.shard], hyper, self.stats,
                          self.comm_status, self.commits, max_ctas=self.UPD_CTAS)

    def _after_update(self, b: int) -> None:
        k = self._k(b)
        if self.resident[k] == b:
            self.resident[k] = -1                      # the staged copy of bucket b is stale now

    # ---- checkpoints: compact shards with their flat ranges ------------------------------------------------------
    def checkpoint_segments(self):
        r, out = self.env.rank, []
        for u in self.units:
            lo, hi = u.off + r * u.shard, u.off + (r + 1) * u.shard
            sl = slice(u.coff, u.coff + u.shard)
            out += [("p", lo, hi, self.p_own[sl]), ("m", lo, hi, self.m_own[sl]), ("v", lo, hi, self.v_own[sl])]
        return out

    def restore_targets(self, n_valid: int):
        return self.checkpoint_segments()

    def load_state(self, ckpt: dict) -> None:
        if "_sharded_dir" not in ckpt:
            # reference-layout file: scatter the full tensors into the shards
            r = self.env.rank
            sd, st = ckpt["model"], ckpt["optimizer"]["state"]
            names = [n for n, _, _, _ in self.flat.layout]
            full = {"p": {}, "m": {}, "v": {}}
            for i, n in enumerate(names):
                full["p"][n] = sd[n]
                full["m"][n], full["v"][n] = st[i]["exp_avg"], st[i]["exp_avg_sq"]
            for u in self.units:
                for key, dst in (("p", self.p_own), ("m", self.m_own), ("v", self.v_own)):
                    if u.name.endswith("attention.wqkv"):
                        pre = u.name[:-len("wqkv")]
                        src = torch.cat([full[key][pre + k + ".weight"].reshape(-1) for k in ("wq", "wk", "wv")])
                    else:
                        src = full[key][u.name].reshape(-1)
                    dst[u.coff:u.coff + u.shard].copy_(src[r * u.shard:(r + 1) * u.shard])
            g = ckpt["optimizer"]["param_groups"][0]
            for k, v in g.items():
                if k != "params":
                    self.optimizer.param_groups[0][k] = v
            self.optimizer.steps_done = int(float(next(iter(st.values()))["step"])) if st else 0
            self.lr_scheduler.load_state_dict(ckpt["lr_scheduler"])
            torch.cuda.synchronize()
            self.env.barrier()
        else:
            super().load_state(ckpt)
        self.resident = [-1] * len(self.resident)

    def full_parameters(self) -> Dict[str, torch.Tensor]:
        """All-gather every parameter (testing / export); not used on the training path."""
        torch.cuda.synchronize()
        self.env.barrier()
        out = {}
        for u in self.units:
            buf = torch.empty(u.numel, dtype=torch.bfloat16, device=self.device)
            N.fsdp_gather(self.symm.table("own"), u.coff, buf, u.shard, self.env.rank, self.env.world_size, self.comm_status)
            out[u.name] = buf.view(u.rows, u.cols) if u.rows > 1 else buf
        torch.cuda.synchronize()
        self.env.barrier()
        return out
