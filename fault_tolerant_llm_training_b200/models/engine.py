"""B200 execution engine: hand-scheduled forward/backward over the sm_100a kernels.

The reference lets autograd drive ~40 library kernels per layer.  Here one training
step is a fixed launch sequence over preallocated buffers:

  * parameters, gradients and both Adam moments are four flat bf16 buffers (`FlatState`); the
    `nn.Parameter`s of the model are views into them, so `state_dict()` / checkpoints are unchanged;
  * wq|wk|wv are adjacent in the flat buffer, so QKV is one [T, d] x [d, 6144] GEMM and its weight
    gradient is one GEMM written straight into the gradient buffer (no autograd accumulation pass);
  * RMSNorm+residual, RoPE, SwiGLU, cross-entropy, grad-norm and AdamW are the fused kernels of
    csrc/ (each replaces 3-10 ATen kernels of the reference);
  * activations are saved in static per-layer buffers (8.7 GB at seq 2048 — 180 GB of HBM makes
    recomputation pointless here), which also makes the step CUDA-graph capturable.

Every GEMM of the step is the hand-written tcgen05 CTA-pair kernel of csrc/gemm_tc.cu (RoPE, SwiGLU forward and
SwiGLU backward in its epilogues, gradient-norm partials out of the weight-gradient epilogues); attention is the
tcgen05 flash kernel of csrc/attn_tc.cu (deterministic forward and backward).  `FTL_GEMM=cublas` / `FTL_ATTN=lib`
switch either back to the library op for A/B comparisons.
"""
from __future__ import annotations

import math
import os
from typing import List, Optional

import torch
import torch.nn.functional as F

from ..config import ModelConfig
from ..ops import native as N
from ..optim.lr import build_lr_scheduler
from ..parallel.env import DistEnv
from .llama import Transformer, build_model, flat_size, param_layout


# --------------------------------------------------------------------------------------------------
class FlatState:
    """Flat bf16 buffers for params / grads / exp_avg / exp_avg_sq with the model's tensors as views."""

    def __init__(self, model: Transformer, device: torch.device, buffers=None, pad_to: int = 128):
        """buffers: optional dict name -> callable(numel) returning a zeroed bf16 tensor for "p" / "g" / "m" / "v"
        (the data-parallel engine puts p and g in peer-mapped memory); pad_to rounds the flat length up so it
        splits evenly into per-rank shards."""
        self.layout = param_layout(model)
        self.numel = (flat_size(self.layout) + pad_to - 1) // pad_to * pad_to
        default = lambda n: torch.zeros(n, dtype=torch.bfloat16, device=device)
        buffers = buffers or {}
        self.p, self.g = buffers.get("p", default)(self.numel), buffers.get("g", default)(self.numel)
        self.m, self.v = buffers.get("m", default)(self.numel), buffers.get("v", default)(self.numel)
        self.index = {}
        params = dict(model.named_parameters())
        with torch.no_grad():
            for name, shape, off, n in self.layout:
                prm = params[name]
                view = self.p[off:off + n].view(shape)
                view.copy_(prm.data)
                prm.data = view
                prm.grad = self.g[off:off + n].view(shape)
                self.index[name] = (off, n, shape)

    # views used by the engine; the fully-sharded state overrides the parameter ones (weights live in staging slots)
    def param_view(self, name: str) -> torch.Tensor:
        return self.view(self.p, name)

    def grad_view(self, name: str) -> torch.Tensor:
        return self.view(self.g, name)

    def param_span(self, first: str, last: str, shape) -> torch.Tensor:
        return self.span(self.p, first, last, shape)

    def grad_span(self, first: str, last: str, shape) -> torch.Tensor:
        return self.span(self.g, first, last, shape)

    def view(self, buf: torch.Tensor, name: str) -> torch.Tensor:
        off, n, shape = self.index[name]
        return buf[off:off + n].view(shape)

    def span(self, buf: torch.Tensor, first: str, last: str, shape) -> torch.Tensor:
        """One view covering adjacent tensors first..last (e.g. wq|wk|wv as [6144, 4096])."""
        o0, _, _ = self.index[first]
        o1, n1, _ = self.index[last]
        v = buf[o0:o1 + n1]
        assert v.numel() == math.prod(shape), "tensors are not adjacent without padding"
        return v.view(shape)


class FlatAdamW(torch.optim.Optimizer):
    """torch.optim.AdamW-compatible front end over the fused flat kernel (csrc/optim.cu).

    `state_dict()` has exactly torch AdamW's layout (per-parameter `step` fp32 0-d CPU tensor, bf16
    `exp_avg` / `exp_avg_sq`; param_groups with the same keys), so the reference's
    `optimizer.load_state_dict(checkpoint["optimizer"])` (train.py:70-72) accepts it and vice versa.
    """

    def __init__(self, model: Transformer, flat: FlatState, lr: float):
        defaults = dict(lr=lr, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, amsgrad=False, maximize=False,
                        foreach=None, capturable=False, differentiable=False, fused=None,
                        decoupled_weight_decay=True)
        super().__init__(list(model.parameters()), defaults)
        self.flat = flat
        self.steps_done = 0
        self._names = [n for n, _ in model.named_parameters()]
        if flat.m.numel():   # the fully-sharded engine keeps only compact moment shards (no per-parameter views)
            for name, prm in zip(self._names, self.param_groups[0]["params"]):
                self.state[prm] = {"step": torch.tensor(0.0, dtype=torch.float32),
                                   "exp_avg": flat.view(flat.m, name), "exp_avg_sq": flat.view(flat.v, name)}

    def hyper_values(self, max_norm: float) -> List[float]:
        """hyper[8] for the step about to be taken (step count t = steps_done + 1)."""
        g = self.param_groups[0]
        b1, b2 = g["betas"]
        t = self.steps_done + 1
        return [g["lr"], b1, b2, g["eps"], g["weight_decay"], 1.0 - b1 ** t, math.sqrt(1.0 - b2 ** t), max_norm]

    def zero_grad(self, set_to_none: bool = False) -> None:  # grads are overwritten, never accumulated
        pass

    @torch.no_grad()
    def step(self, hyper_dev: torch.Tensor = None, stats_dev: torch.Tensor = None, closure=None):
        f = self.flat
        N.adamw_step(f.p, f.g, f.m, f.v, hyper_dev, stats_dev)
        self.steps_done += 1

    def state_dict(self):
        for st in self.state.values():
            st["step"].fill_(float(self.steps_done))
        return super().state_dict()

    @torch.no_grad()
    def load_state_dict(self, state_dict) -> None:
        groups = state_dict["param_groups"]
        assert len(groups) == 1 and len(groups[0]["params"]) == len(self._names), "parameter count mismatch"
        for k, v in groups[0].items():
            if k != "params":
                self.param_groups[0][k] = v
        steps = 0
        for idx, name in zip(groups[0]["params"], self._names):
            st = state_dict["state"].get(idx)
            if st is None:
                continue
            self.flat.view(self.flat.m, name).copy_(st["exp_avg"])
            self.flat.view(self.flat.v, name).copy_(st["exp_avg_sq"])
            steps = max(steps, int(float(st["step"])))
        self.steps_done = steps


# --------------------------------------------------------------------------------------------------
class LossHandle:
    """Loss of one step: a pinned host slot + the event recorded after its device->host copy."""

    def __init__(self, host: torch.Tensor, event: torch.cuda.Event):
        self.host, self.event = host, event

    def item(self) -> float:
        self.event.synchronize()
        return float(self.host[0])

    def grad_norm(self) -> float:
        self.event.synchronize()
        return float(self.host[1])

    def finite(self) -> bool:
        self.event.synchronize()
        return bool(self.host[3] != 0.0)

    def __float__(self) -> float:
        return self.item()


class _Nvtx:
    """NVTX ranges per phase and layer (`FTL_NVTX=1`; off by default: ~300 push/pop pairs per step cost host time)."""
    on = os.environ.get("FTL_NVTX", "0") == "1"

    @classmethod
    def push(cls, name: str) -> None:
        if cls.on:
            torch.cuda.nvtx.range_push(name)

    @classmethod
    def pop(cls) -> None:
        if cls.on:
            torch.cuda.nvtx.range_pop()


class NativeEngine:
    """Forward + backward of the decoder for a fixed (batch, seq) shape."""

    def __init__(self, model: Transformer, flat: FlatState, cfg: ModelConfig, batch: int, seq: int,
                 device: torch.device):
        self.model, self.flat, self.cfg = model, flat, cfg
        self.B, self.S, self.T = batch, seq, batch * seq
        self.dev = device
        c, T = cfg, self.T
        d, Fh, H, KV, hd, V = c.dim, c.ffn_hidden, c.n_heads, c.n_kv_heads, c.head_dim, c.vocab_size
        self.qkv_dim = c.qkv_dim
        bf = lambda *s: torch.empty(*s, dtype=torch.bfloat16, device=device)
        f32 = lambda *s: torch.empty(*s, dtype=torch.float32, device=device)
        L = c.n_layers
        # saved activations, one set per layer
        self.x = [bf(T, d) for _ in range(L + 1)]       # block inputs; x[L] = final residual stream
        self.xn = [bf(T, d) for _ in range(L)]
        self.rstd1 = [f32(T) for _ in range(L)]
        self.qkv = [bf(T, self.qkv_dim) for _ in range(L)]
        self.attn = [None] * L                           # attention outputs (graph-holding tensors)
        self.attn_in = [None] * L
        self.own_attn = self.ATTN_TC and seq % 128 == 0 and hd in (64, 128)
        if self.own_attn:
            self.attn_o = [bf(T, H * hd) for _ in range(L)]     # saved for the backward (delta, wo weight gradient)
            self.attn_lse = [f32(batch * H * seq) for _ in range(L)]
            self.attn_delta = f32(batch * H * seq)
        self.ao = bf(T, d)                               # wo output (consumed immediately)
        self.h = [bf(T, d) for _ in range(L)]
        self.hn = [bf(T, d) for _ in range(L)]
        self.rstd2 = [f32(T) for _ in range(L)]
        self.gate = [bf(T, Fh) for _ in range(L)]
        self.up = [bf(T, Fh) for _ in range(L)]
        self.act = [bf(T, Fh) for _ in range(L)]
        self.ff = bf(T, d)
        self.xf = bf(T, d)                               # final normed stream
        self.rstd_f = f32(T)
        # FTL_FUSED_HEAD=1: vocabulary-chunked LM head + cross-entropy, no [T, V] buffer (ops/fused_head.py)
        self.head = None
        if self.USE_TC and os.environ.get("FTL_FUSED_HEAD", "0") == "1":
            from ..ops.fused_head import ChunkedHead
            self.head = ChunkedHead(T, d, V, device, chunk=int(os.environ.get("FTL_HEAD_CHUNK", "8192")))
        self.logits = bf(T, V) if self.head is None else None
        self.row_loss = f32(T)
        self.loss = f32(1)
        # backward transients
        self.dx_a, self.dx_b = bf(T, d), bf(T, d)
        self.dh = bf(T, d)
        self.dtmp = bf(T, d)
        # SwiGLU backward rides in the epilogue of the w2 dgrad GEMM (no dact buffer) when the 32-byte row accesses fit
        self.fused_swiglu_bwd = self.USE_TC and Fh % 16 == 0
        self.dact = None if self.fused_swiglu_bwd else bf(T, Fh)
        self.dgate, self.dup = bf(T, Fh), bf(T, Fh)
        self.dqkv = bf(T, self.qkv_dim)
        self.dattn = bf(T, H * hd)
        self.norm_scratch = f32(N.rmsnorm_partials(T) * d)
        # rope tables (fp32 [seq, hd/2])
        self.cos = model.rope_cos[:seq].to(device=device, dtype=torch.float32).contiguous()
        self.sin = model.rope_sin[:seq].to(device=device, dtype=torch.float32).contiguous()
        # weight views
        pv, gv = flat.param_view, flat.grad_view
        self.w_emb, self.g_emb = pv("tok_embeddings.weight"), gv("tok_embeddings.weight")
        self.w_out, self.g_out = pv("output.weight"), gv("output.weight")
        self.w_fn, self.g_fn = pv("norm.weight"), gv("norm.weight")
        self.lw = []
        for i in range(L):
            b = f"layers.{i}."
            names = dict(q=b + "attention.wq.weight", v=b + "attention.wv.weight", o=b + "attention.wo.weight",
                         w1=b + "feed_forward.w1.weight", w2=b + "feed_forward.w2.weight",
                         w3=b + "feed_forward.w3.weight", an=b + "attention_norm.weight", fn=b + "ffn_norm.weight")
            lw = {"wqkv": flat.param_span(names["q"], names["v"], (self.qkv_dim, d)),
                  "gqkv": flat.grad_span(names["q"], names["v"], (self.qkv_dim, d))}
            for k in ("o", "w1", "w2", "w3", "an", "fn"):
                lw["w" + k], lw["g" + k] = pv(names[k]), gv(names[k])
            self.lw.append(lw)
        # Pipelining hooks (parallel/pipeline.py).  Buckets: 0 = embedding, 1+i = layer i, L+1 = final norm + head.
        #   hook_params(b): called before the first kernel that reads bucket b's parameters
        #   hook_grads(b):  called after the last kernel that writes bucket b's gradients
        self.hook_params = None
        self.hook_grads = None
        self.sparse_emb_zero, self._emb_dirty = False, None   # see backward(): set by single-rank trainers
        # Gradient-norm partials produced by the kernels that write the gradients (single-rank pipelined trainer): one
        # slice per producing launch, handed out in launch order, summed in index order by the publish kernel.
        self.norm_arena: Optional[torch.Tensor] = None
        self._arena_off = 0
        self.use_graphs = (os.environ.get("FTL_GRAPHS", "0") == "1" and self.USE_TC and self.own_attn)
        self._graphs, self._steps_seen = {}, 0
        self.hook_params_bwd = None   # called before the backward of bucket b reads its parameters
        self.hook_tokens = None       # called at the start of the backward with the step's flat token ids
        self.fused_gather = None      # FSDP: object with qkv(i, x, out) / swiglu(i, x, act, gate, up) running the
                                      # gather-fused GEMMs for layer i (parallel/fsdp.py)

    def enable_fused_norm(self) -> bool:
        """Allocate the arena for sum-of-squares partials written by the weight-gradient GEMMs, the norm-weight
        finalize kernels and the embedding backward.  Replaces a 16 GB re-read of the gradient buffer per step."""
        if not self.USE_TC:
            return False
        c = self.cfg
        gs, ns = N.gemm_sumsq_slots(), N.rmsnorm_sumsq_slots(c.dim)
        head_gemms = self.head.n_chunks if self.head is not None else 1     # one weight-gradient GEMM per vocabulary chunk
        n = c.n_layers * (5 * gs + 2 * ns) + head_gemms * gs + ns + self.T
        self.norm_arena = torch.zeros(n, dtype=torch.float32, device=self.dev)
        return True

    def _sq(self, n: int) -> Optional[torch.Tensor]:
        if self.norm_arena is None:
            return None
        o = self._arena_off
        self._arena_off = o + n
        assert self._arena_off <= self.norm_arena.numel()
        return self.norm_arena[o:o + n]

    def _sq_gemm(self):
        return self._sq(N.gemm_sumsq_slots()) if self.norm_arena is not None else None

    def _sq_norm(self):
        return self._sq(N.rmsnorm_sumsq_slots(self.cfg.dim)) if self.norm_arena is not None else None

    def _need(self, upto: int) -> None:
        if self.hook_params is not None:
            while self._params_waited < upto:
                self._params_waited += 1
                self.hook_params(self._params_waited)

    # ---- GEMM dispatch: C = A @ B^T (nt), A @ B (nn), A^T @ B (tn) -------------------------------
    # FTL_GEMM=tc (default): the hand-written tcgen05 CTA-pair kernel of csrc/gemm_tc.cu for every GEMM of the
    # step, with the RoPE / SwiGLU epilogues and the split reduction fused (0.97-1.20x cuBLAS on these shapes,
    # profiles/gemm_tc_pair_vs_cublas.jsonl).  FTL_GEMM=cublas: library GEMMs + separate elementwise kernels.
    USE_TC = os.environ.get("FTL_GEMM", "tc") != "cublas"
    # FTL_ATTN=tc: the hand-written tcgen05 flash attention of csrc/attn_tc.cu (deterministic forward and backward,
    # dq|dk|dv written in place with the inverse RoPE fused).  FTL_ATTN=lib: library SDPA through a small autograd island.
    ATTN_TC = os.environ.get("FTL_ATTN", "tc") != "lib"

    def gemm_nt(self, a, b, out):
        if self.USE_TC:
            return N.gemm(a, b, out)
        return torch.mm(a, b.t(), out=out)

    def gemm_nn(self, a, b, out, a2=None, b2=None):
        """out = a @ b (+ a2 @ b2)."""
        if self.USE_TC:
            return N.gemm(a, b, out, b_mn=True, a2=a2, b2=b2)
        torch.mm(a, b, out=out)
        if a2 is not None:
            torch.addmm(out, a2, b2, out=out)
        return out

    def gemm_tn(self, a, b, out):
        """Weight gradient out = a^T @ b; also the producer of its share of the gradient-norm partials."""
        if self.USE_TC:
            return N.gemm(a, b, out, a_mn=True, b_mn=True, sumsq=self._sq_gemm())
        return torch.mm(a.t(), b, out=out)

    # ---- attention (library SDPA island; replaced by csrc/attn when available) -------------------
    def _attention_fwd(self, i: int) -> torch.Tensor:
        c, B, S = self.cfg, self.B, self.S
        H, KV, hd = c.n_heads, c.n_kv_heads, c.head_dim
        qkv = self.qkv[i]
        if self.own_attn:
            return N.attn_fwd(qkv, self.attn_o[i], self.attn_lse[i], B, S, H, KV, hd)
        q = qkv[:, :H * hd].view(B, S, H, hd).transpose(1, 2).detach().requires_grad_()
        k = qkv[:, H * hd:(H + KV) * hd].view(B, S, KV, hd).transpose(1, 2).detach().requires_grad_()
        v = qkv[:, (H + KV) * hd:].view(B, S, KV, hd).transpose(1, 2).detach().requires_grad_()
        with torch.enable_grad():
            o = F.scaled_dot_product_attention(q, k, v, is_causal=True, enable_gqa=(H != KV))
        self.attn[i], self.attn_in[i] = o, (q, k, v)
        return o.transpose(1, 2).reshape(self.T, H * hd)

    def _attention_bwd(self, i: int, dout: torch.Tensor) -> None:
        c, B, S = self.cfg, self.B, self.S
        H, KV, hd = c.n_heads, c.n_kv_heads, c.head_dim
        if self.own_attn:   # dq|dk|dv land in self.dqkv with the inverse rotary embedding already applied
            N.attn_bwd(self.qkv[i], self.attn_o[i], dout, self.attn_lse[i], self.attn_delta, self.dqkv, B, S, H, KV, hd,
                       self.cos, self.sin)
            return
        o, (q, k, v) = self.attn[i], self.attn_in[i]
        do = dout.view(B, S, H, hd).transpose(1, 2)
        dq, dk, dv = torch.autograd.grad(o, (q, k, v), do)
        d = self.dqkv
        d[:, :H * hd].view(B, S, H, hd).copy_(dq.transpose(1, 2))
        d[:, H * hd:(H + KV) * hd].view(B, S, KV, hd).copy_(dk.transpose(1, 2))
        d[:, (H + KV) * hd:].view(B, S, KV, hd).copy_(dv.transpose(1, 2))
        self.attn[i] = self.attn_in[i] = None

    # ---- per-layer CUDA graphs (FTL_GRAPHS=1) ---------------------------------------------------------------------
    # A block's forward is 7 launches and its backward 13, all over static buffers: after two eager steps (so that every
    # one-time attribute call and scratch allocation has happened) each is captured once and replayed, which removes
    # the 2-5 us launch gaps inside a block (~1.9 ms of a step's compute stream, profiles/step_gaps_r1.md).  Everything
    # that takes a host-side decision stays outside the graphs: parameter / gradient hooks (event waits, reduce-scatter
    # launches), the head, the embedding.
    def _layer(self, kind: str, i: int, fn, *args) -> None:
        # (the fully-sharded engine keeps host-side bookkeeping inside its gather-fused GEMM calls: no graphs there)
        if not self.use_graphs or self.fused_gather is not None or self.hook_params_bwd is not None:
            return fn(*args)
        key = (kind, i)
        g = self._graphs.get(key)
        if g is not None:
            g[0].replay()
            self._arena_off = g[1]
            N.KERNELS += g[2]                             # the kernels inside the graph ran: keep the launch count honest
            N.LAUNCHES += 1
            return
        if self._steps_seen < 2:
            return fn(*args)
        graph, k0 = torch.cuda.CUDAGraph(), N.KERNELS
        with torch.cuda.graph(graph, capture_error_mode="thread_local"):
            fn(*args)
        # the partial-sum slices taken inside are baked into the graph; remember where the bump allocator ended
        self._graphs[key] = (graph, self._arena_off, N.KERNELS - k0)
        graph.replay()                                    # capture only records: run the block for this step

    def _fwd_layer(self, i: int) -> None:
        """The seven launches of block i's forward (no host-side decisions: capturable as one CUDA graph)."""
        c, lw = self.cfg, self.lw[i]
        H, KV, hd = c.n_heads, c.n_kv_heads, c.head_dim
        rot = (H + KV) * hd
        # residual add of the previous block fused with this block's attention_norm (x_i = h_{i-1} + ff)
        if i == 0:
            N.rmsnorm_fwd(self.x[0], lw["wan"], c.norm_eps, self.xn[0], self.rstd1[0])
        else:
            N.rmsnorm_fwd(self.h[i - 1], lw["wan"], c.norm_eps, self.xn[i], self.rstd1[i], delta=self.ff, h_out=self.x[i])
        if self.fused_gather is not None:   # FSDP: wq|wk|wv arrive over NVLink inside the GEMM (N12)
            self.fused_gather.qkv(i, self.xn[i], self.qkv[i])
        elif self.USE_TC:   # QKV projection with RoPE in the GEMM epilogue (K3 + K4)
            N.gemm_qkv_rope(self.xn[i], lw["wqkv"], self.qkv[i], self.cos, self.sin, self.S, hd, rot)
        else:
            self.gemm_nt(self.xn[i], lw["wqkv"], self.qkv[i])
            N.rope_inplace(self.qkv[i], self.cos, self.sin, self.S, rot, hd)
        a = self._attention_fwd(i)
        self.gemm_nt(a, lw["wo"], self.ao)
        # h = x + attn_out ; hn = norm(h)
        N.rmsnorm_fwd(self.x[i], lw["wfn"], c.norm_eps, self.hn[i], self.rstd2[i], delta=self.ao, h_out=self.h[i])
        if self.fused_gather is not None:
            self.fused_gather.swiglu(i, self.hn[i], self.act[i], self.gate[i], self.up[i])
        elif self.USE_TC:   # gate|up dual GEMM with the SwiGLU epilogue (K9 + K10)
            N.gemm_swiglu(self.hn[i], lw["ww1"], lw["ww3"], self.act[i], self.gate[i], self.up[i])
        else:
            self.gemm_nt(self.hn[i], lw["ww1"], self.gate[i])
            self.gemm_nt(self.hn[i], lw["ww3"], self.up[i])
            N.swiglu_fwd(self.gate[i], self.up[i], self.act[i])
        self.gemm_nt(self.act[i], lw["ww2"], self.ff)

    def _bwd_layer(self, i: int, dx: torch.Tensor, spare: torch.Tensor) -> None:
        """The launches of block i's backward: d(x_{i+1}) in `dx` -> d(x_i) in `spare`, weight gradients in place."""
        c, T, lw = self.cfg, self.T, self.lw[i]
        H, KV, hd = c.n_heads, c.n_kv_heads, c.head_dim
        rot = (H + KV) * hd
        # x_{i+1} = h + w2(act)
        self.gemm_tn(dx, self.act[i], lw["gw2"])
        if self.fused_swiglu_bwd:   # d act -> (d gate, d up) in the dgrad epilogue
            N.gemm_swiglu_bwd(dx, lw["ww2"], self.gate[i], self.up[i], self.dgate, self.dup)
        else:
            self.gemm_nn(dx, lw["ww2"], self.dact)
            N.swiglu_bwd(self.dact, self.gate[i], self.up[i], self.dgate, self.dup)
        self.gemm_tn(self.dgate, self.hn[i], lw["gw1"])
        self.gemm_tn(self.dup, self.hn[i], lw["gw3"])
        self.gemm_nn(self.dgate, lw["ww1"], self.dtmp, a2=self.dup, b2=lw["ww3"])   # one pass over K = 2F
        # dh = dx + d ffn_norm
        N.rmsnorm_bwd(self.dtmp, self.h[i], lw["wfn"], self.rstd2[i], self.dh, lw["gfn"], self.norm_scratch,
                      dres=dx, sumsq=self._sq_norm())
        # h = x + wo(attn)
        a = self.attn_o[i] if self.own_attn else self.attn[i].transpose(1, 2).reshape(T, H * hd)
        self.gemm_tn(self.dh, a, lw["go"])
        self.gemm_nn(self.dh, lw["wo"], self.dattn)
        self._attention_bwd(i, self.dattn)
        if not self.own_attn:
            N.rope_inplace(self.dqkv, self.cos, self.sin, self.S, rot, hd, inverse=True)
        self.gemm_tn(self.dqkv, self.xn[i], lw["gqkv"])
        self.gemm_nn(self.dqkv, lw["wqkv"], self.dtmp)
        N.rmsnorm_bwd(self.dtmp, self.x[i], lw["wan"], self.rstd1[i], spare, lw["gan"], self.norm_scratch,
                      dres=self.dh, sumsq=self._sq_norm())

    # ---- forward ---------------------------------------------------------------------------------
    @torch.no_grad()
    def forward(self, tokens: torch.Tensor, labels: torch.Tensor, inv_n: torch.Tensor, write_grad: bool = True):
        c, T = self.cfg, self.T
        H, KV, hd = c.n_heads, c.n_kv_heads, c.head_dim
        rot = (H + KV) * hd
        self._params_waited = -1
        self._need(0)
        _Nvtx.push("forward")
        N.embedding_fwd(tokens.view(-1), self.w_emb, self.x[0])
        for i, lw in enumerate(self.lw):
            if i:
                _Nvtx.pop()
            _Nvtx.push(f"fwd layer {i}")
            self._need(i + 1)
            self._layer("f", i, self._fwd_layer, i)
        L = c.n_layers
        _Nvtx.pop()
        _Nvtx.push("fwd head + loss")
        self._need(L + 1)
        # x_L = h_{L-1} + ff, fused with the final norm
        N.rmsnorm_fwd(self.h[L - 1], self.w_fn, c.norm_eps, self.xf, self.rstd_f, delta=self.ff, h_out=self.x[L])
        if self.head is not None:
            self.head.forward(self.xf, self.w_out, labels.view(-1), inv_n, self.row_loss, self.loss)
            self._head_args = (labels.view(-1), inv_n)      # the backward recomputes the logits chunk by chunk
        else:
            self.gemm_nt(self.xf, self.w_out, self.logits)
            N.cross_entropy(self.logits, labels.view(-1), inv_n, self.row_loss, self.loss, write_grad=write_grad)
        _Nvtx.pop()
        _Nvtx.pop()
        return self.loss

    # ---- backward (logits already hold d loss / d logits) ---------------------------------------------
    @torch.no_grad()
    def backward(self, tokens: torch.Tensor):
        c, T = self.cfg, self.T
        H, KV, hd = c.n_heads, c.n_kv_heads, c.head_dim
        rot = (H + KV) * hd
        L = c.n_layers
        dl = self.logits
        self._arena_off = 0
        if self.hook_tokens is not None:
            self.hook_tokens(tokens.view(-1))
        _Nvtx.push("backward")
        _Nvtx.push("bwd head")
        if self.hook_params_bwd is not None:
            self.hook_params_bwd(L + 1)
        if self.head is not None:
            labels, inv_n = self._head_args
            self.head.backward(self.xf, self.w_out, labels, inv_n, self.g_out, self.dtmp, sumsq=self._sq_gemm)
        else:
            self.gemm_tn(dl, self.xf, self.g_out)                     # dW_out
            self.gemm_nn(dl, self.w_out, self.dtmp)                   # d xf
        dx, spare = self.dx_a, self.dx_b
        N.rmsnorm_bwd(self.dtmp, self.x[L], self.w_fn, self.rstd_f, dx, self.g_fn, self.norm_scratch, sumsq=self._sq_norm())
        if self.hook_grads is not None:
            self.hook_grads(L + 1)
        for i in range(L - 1, -1, -1):
            lw = self.lw[i]
            _Nvtx.pop()
            _Nvtx.push(f"bwd layer {i}")
            if self.hook_params_bwd is not None:
                self.hook_params_bwd(i + 1)
            self._layer("b", i, self._bwd_layer, i, dx, spare)
            dx, spare = spare, dx
            if self.hook_grads is not None:
                self.hook_grads(i + 1)
        # embedding: deterministic scatter-add into a zero table gradient, without a sort (csrc/elementwise.cu).  The
        # table gradient is 1 GB and only the rows of this batch's tokens are written, so when nobody else writes into
        # it (single rank: no in-place reduction of peers' rows) the kernel itself clears the rows the PREVIOUS step
        # touched and this one does not, instead of a memset of the whole buffer.
        _Nvtx.pop()
        _Nvtx.push("bwd embedding")
        flat_tok = tokens.view(-1)
        prev = None
        if self.sparse_emb_zero and self._emb_dirty is not None:
            prev = self._emb_dirty
        else:
            self.g_emb.zero_()
        N.embedding_bwd_direct(flat_tok, prev, dx, self.g_emb, sumsq=self._sq(self.T))
        if self.sparse_emb_zero:
            if self._emb_dirty is None:
                self._emb_dirty = torch.empty_like(flat_tok)
            self._emb_dirty.copy_(flat_tok)   # device-to-device copy; the kernel ignores ids outside the table
        if self.hook_grads is not None:
            self.hook_grads(0)
        _Nvtx.pop()
        _Nvtx.pop()
        self._steps_seen += 1


# --------------------------------------------------------------------------------------------------
class NativeTrainer:
    """`.step(inputs, labels)` on a B200: staged H2D, fwd, bwd, grad norm (+clip), fused AdamW, LR step."""

    RING = 4

    def __init__(self, cfg: ModelConfig, env: DistEnv, *, dtype: torch.dtype, lr: float, warmup_steps: int,
                 grad_max_norm: float, clip_mode: str = "clip", seed: int = 0, parallel: str = "none",
                 keep_shards: bool = False,
                 consolidate: bool = True):
        if dtype != torch.bfloat16:
            raise ValueError("the B200 engine is bf16-only (use --device cpu for other dtypes)")
        N.lib()  # fail loudly if the sm_100a library is missing
        # Bitwise-reproducible steps are part of the contract (kill/resume chains must replay exactly).  Everything in
        # csrc/ is deterministic by construction (fixed reduction orders, no float atomics); the one library op left
        # on the path, SDPA's backward, accumulates dQ with atomics at long sequence lengths unless told otherwise.
        if os.environ.get("FTL_ATTN", "tc") == "lib" and os.environ.get("FTL_DETERMINISTIC", "1") == "1":
            torch.use_deterministic_algorithms(True, warn_only=True)   # only the library attention path needs it
        self.cfg, self.env, self.device = cfg, env, env.device
        self.parallel = parallel
        self.grad_max_norm, self.clip_mode = float(grad_max_norm), clip_mode
        self.model = build_model(cfg, dtype, self.device, seed=seed)
        self.model.train()
        self.flat = self._make_flat_state()
        self.optimizer = FlatAdamW(self.model, self.flat, lr)
        self.lr_scheduler = build_lr_scheduler(self.optimizer, warmup_steps)
        self.engine: Optional[NativeEngine] = None
        self.hyper = torch.zeros(8, dtype=torch.float32, device=self.device)
        self.stats = torch.zeros(4, dtype=torch.float32, device=self.device)
        self.gn_scratch = torch.zeros(N.grad_norm_partials(), dtype=torch.float32, device=self.device)
        self.result_dev = torch.zeros(8, dtype=torch.float32, device=self.device)
        self.result_host = [torch.zeros(8, dtype=torch.float32).pin_memory() for _ in range(self.RING)]
        self.events = [torch.cuda.Event() for _ in range(self.RING)]
        self._slot = 0
        self._stage = None
        self._pending: List[LossHandle] = []   # results not yet inspected by the host (at most 2)
        self.last_grad_norm = 0.0
        self.h2d_bytes_per_step = 0
        self.d2h_bytes_per_step = 32
        self.kernel_launches_per_step = 0
        from ..ckpt.sharded import FlatCheckpointer
        self.checkpointer = FlatCheckpointer(self, keep_shards=keep_shards, consolidate=consolidate)

    def _make_flat_state(self) -> FlatState:
        return FlatState(self.model, self.device)

    # -- per-shape setup --------------------------------------------------------------------------
    def _ensure_engine(self, B: int, S: int) -> None:
        if self.engine is not None and (self.engine.B, self.engine.S) == (B, S):
            return
        self.engine = NativeEngine(self.model, self.flat, self.cfg, B, S, self.device)
        self.engine.sparse_emb_zero = self.env.world_size == 1 and self.parallel != "fsdp"
        if self.engine.sparse_emb_zero:
            self.engine.g_emb.zero_()                          # a previous engine (other shape) may have left rows behind
        T = B * S
        nbytes = 2 * T * 8 + 64
        # staging ring: [tokens | labels | inv_n, hyper[8]] packed so that a step needs ONE H2D copy.  The host
        # runs at most two steps ahead of the GPU (see step()), so a ring of RING slots is never overwritten
        # while its copy is in flight.
        self._stage = []
        for _ in range(self.RING):
            host = torch.zeros(nbytes, dtype=torch.uint8).pin_memory()
            dev = torch.zeros(nbytes, dtype=torch.uint8, device=self.device)
            self._stage.append(dict(
                host=host, dev=dev,
                tok_h=host[:T * 8].view(torch.int64).view(B, S), lab_h=host[T * 8:2 * T * 8].view(torch.int64).view(B, S),
                flt_h=host[2 * T * 8:].view(torch.float32),
                tok_d=dev[:T * 8].view(torch.int64).view(B, S), lab_d=dev[T * 8:2 * T * 8].view(torch.int64).view(B, S),
                flt_d=dev[2 * T * 8:].view(torch.float32)))
        self.h2d_bytes_per_step = nbytes

    def _global_token_count(self, n_local: int) -> int:
        return self.env.sum_int(n_local)

    # -- one step (reference: train.py:93-110) ----------------------------------------------------
    def step(self, inputs: torch.Tensor, labels: torch.Tensor, n_items: Optional[int] = None) -> LossHandle:
        B, S = inputs.shape
        self._ensure_engine(B, S)
        launches0 = N.LAUNCHES
        # results are inspected two steps late, so the host never drains the GPU queue; a non-finite
        # gradient norm made that step's update a no-op on the device (csrc/optim.cu), so raising late
        # still checkpoints clean state (reference: get_total_norm(error_if_nonfinite=True), utils.py:61)
        while len(self._pending) >= 2:
            self._inspect(self._pending.pop(0))
        if n_items is None:
            n_items = self._global_token_count(int(labels.ne(-100).sum()))
        slot = self._slot
        self._slot = (slot + 1) % self.RING
        st = self._stage[slot]
        st["tok_h"].copy_(inputs)
        st["lab_h"].copy_(labels)
        flt = [1.0 / max(n_items, 1)] + self.optimizer.hyper_values(self.grad_max_norm)
        st["flt_h"][:len(flt)] = torch.tensor(flt, dtype=torch.float32)
        st["dev"].copy_(st["host"], non_blocking=True)                  # the one H2D copy of the step
        inv_n, hyper = st["flt_d"][0:1], st["flt_d"][1:9]
        self._begin_step(hyper)
        loss = self.engine.forward(st["tok_d"], st["lab_d"], inv_n)
        self.engine.backward(st["tok_d"])
        self._reduce_and_update(hyper, loss)
        self.lr_scheduler.step()
        handle = self._publish_result(loss, slot)
        self._pending.append(handle)
        self.kernel_launches_per_step = N.LAUNCHES - launches0
        return handle

    def _begin_step(self, hyper: torch.Tensor) -> None:
        pass

    def _publish_result(self, loss: torch.Tensor, slot: int) -> "LossHandle":
        """One D2H copy per step: loss, grad norm, clip coefficient, finite flag, comm status."""
        self.result_dev[0:1].copy_(loss)
        self.result_dev[1:4].copy_(self.stats[0:3])
        self._write_status(self.result_dev[4:5])
        self.result_host[slot].copy_(self.result_dev, non_blocking=True)
        self.events[slot].record()
        return LossHandle(self.result_host[slot], self.events[slot])

    def _inspect(self, old: "LossHandle") -> None:
        """Host-side check of a finished step (two steps late, see step())."""
        old.event.synchronize()
        if old.host[4] not in (0.0, 3.0):   # 3 = non-finite norm, reported below with the reference's message
            from ..ft.signals import PeerAbort
            raise PeerAbort(f"collective step aborted (status {int(old.host[4])}: 1 = barrier timeout, 2 = peer abort)")
        if not old.finite() or old.host[4] == 3.0:
            raise FloatingPointError("The total norm of gradients is non-finite, so it cannot be clipped.")
        self.last_grad_norm = old.grad_norm()

    def _write_status(self, dst: torch.Tensor) -> None:
        dst.zero_()

    def drain(self) -> None:
        """Inspect every outstanding step result (called before a checkpoint is taken)."""
        while self._pending:
            self._inspect(self._pending.pop(0))

    def _reduce_and_update(self, hyper: torch.Tensor, loss: torch.Tensor) -> None:
        """Single-GPU: grad norm over the flat gradient, then the fused AdamW pass."""
        f = self.flat
        N.grad_norm(f.g, hyper, self.gn_scratch, self.stats, clip=(self.clip_mode == "clip"))
        self.optimizer.step(hyper, self.stats)

    # -- checkpoint plumbing ----------------------------------------------------------------------
    def load_state(self, ckpt: dict) -> None:
        if "_sharded_dir" in ckpt:
            self._load_sharded(ckpt)
        else:
            self.model.load_state_dict(ckpt["model"], strict=True)  # copies into the flat views
            self.optimizer.load_state_dict(ckpt["optimizer"])
        self.lr_scheduler.load_state_dict(ckpt["lr_scheduler"])
        torch.cuda.synchronize()

    def _moment_ranges(self):
        """Flat ranges of exp_avg / exp_avg_sq this rank maintains."""
        return [(0, self.flat.numel)]

    def checkpoint_segments(self):
        """(buffer, flat_lo, flat_hi, device tensor) pieces this rank writes into its shard file."""
        f, out = self.flat, []
        for lo, hi in self._moment_ranges():
            out += [("p", lo, hi, f.p[lo:hi]), ("m", lo, hi, f.m[lo:hi]), ("v", lo, hi, f.v[lo:hi])]
        return out

    def restore_targets(self, n_valid: int):
        """(buffer, flat_lo, flat_hi, device tensor) pieces this rank needs back from a checkpoint."""
        f = self.flat
        out = [("p", 0, n_valid, f.p[:n_valid])]
        for lo, hi in self._moment_ranges():
            hi = min(hi, n_valid)
            if lo < hi:
                out += [("m", lo, hi, f.m[lo:hi]), ("v", lo, hi, f.v[lo:hi])]
        return out

    def _load_sharded(self, ckpt: dict) -> None:
        """Shard files -> device, straight through the native reader (no CPU state_dict is ever built)."""
        from ..ckpt.sharded import read_range_to_device
        d, m, f = ckpt["_sharded_dir"], ckpt["_manifest"], self.flat
        if [list(x[1]) for x in m["layout"]] != [list(s) for _, s, _, _ in f.layout]:
            raise ValueError("checkpoint was written for a different model shape")
        n = min(m["flat_numel"], f.numel)   # padding at the tail may differ with the world size; the payload does not
        for buf, lo, hi, dst in self.restore_targets(n):
            read_range_to_device(d, m, buf, lo, hi, dst)
        g = m["optimizer"]["param_groups"][0]
        for k, v in g.items():
            self.optimizer.param_groups[0][k] = tuple(v) if k == "betas" else v
        self.optimizer.steps_done = int(m["optimizer"]["steps_done"])

    def save_fn(self):
        return self.checkpointer.save

    def finalize_fn(self):
        return self.checkpointer.finalize

    def loss_value(self, loss: LossHandle) -> float:
        v = loss.item()
        self.last_grad_norm = loss.grad_norm()
        return v

    def close(self, release: bool = False) -> None:
        pass
