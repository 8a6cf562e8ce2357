"""Plain-PyTorch statements of every op of the decoder.

Two jobs: (1) the CPU execution path (BASELINE.json config 1 runs the tiny model
on CPU), (2) the fp32 oracle that each sm_100a kernel is tested against.  Every
function cites the reference lines whose arithmetic (including the rounding
points) it restates.
"""
from __future__ import annotations

import math
from typing import Tuple

import torch
import torch.nn.functional as F


def rmsnorm(x: torch.Tensor, weight: torch.Tensor, eps: float) -> torch.Tensor:
    """x * rsqrt(mean(x^2) + eps) in fp32, rounded to x.dtype, then times weight
    (reference: model.py:43-48 — note the cast happens *before* the weight multiply)."""
    xf = x.float()
    inv = torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)
    return (xf * inv).type_as(x) * weight


def rope_table(head_dim: int, seq_len: int, theta: float, device=None) -> Tuple[torch.Tensor, torch.Tensor]:
    """cos/sin tables [seq_len, head_dim/2] in fp32.  Same angles as the complex64
    table of the reference (model.py:51-71): angle[t, i] = t * theta^(-2i/head_dim)."""
    inv_freq = 1.0 / (theta ** (torch.arange(0, head_dim, 2, device=device)[: head_dim // 2].float() / head_dim))
    t = torch.arange(seq_len, device=device)
    ang = torch.outer(t, inv_freq).float()
    return torch.cos(ang), torch.sin(ang)


def rope_apply(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor, inverse: bool = False) -> torch.Tensor:
    """Rotate adjacent pairs (x[2i], x[2i+1]) by angle[t, i] in fp32, round to
    x.dtype (reference: model.py:121-126, complex multiply on view_as_complex).
    x: [b, s, heads, head_dim].  inverse=True rotates by -angle (the backward)."""
    b, s, h, d = x.shape
    xf = x.float().reshape(b, s, h, d // 2, 2)
    xe, xo = xf[..., 0], xf[..., 1]
    c = cos[:s].view(1, s, 1, d // 2)
    sn = sin[:s].view(1, s, 1, d // 2)
    if inverse:
        sn = -sn
    oe = xe * c - xo * sn
    oo = xe * sn + xo * c
    return torch.stack((oe, oo), dim=-1).reshape(b, s, h, d).type_as(x)


def attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor) -> torch.Tensor:
    """Causal grouped-query attention.  q: [b, s, H, d]; k, v: [b, s, KV, d].
    Returns [b, s, H, d].  The reference materialises repeat_kv (model.py:129-138,
    204-205) and calls SDPA with is_causal=True (model.py:212); the result is the
    same as attending each query head to KV head (h // n_rep)."""
    H, KV = q.shape[2], k.shape[2]
    qt, kt, vt = q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2)
    if H != KV:
        rep = H // KV
        kt = kt.repeat_interleave(rep, dim=1)
        vt = vt.repeat_interleave(rep, dim=1)
    out = F.scaled_dot_product_attention(qt, kt, vt, is_causal=True)
    return out.transpose(1, 2)


def attention_fp32_oracle(q, k, v):
    """Explicit softmax(QK^T/sqrt(d) + causal) V in fp32 — test oracle."""
    b, s, H, d = q.shape
    KV = k.shape[2]
    rep = H // KV
    qf = q.float().transpose(1, 2)
    kf = k.float().transpose(1, 2).repeat_interleave(rep, dim=1)
    vf = v.float().transpose(1, 2).repeat_interleave(rep, dim=1)
    sc = qf @ kf.transpose(-1, -2) / math.sqrt(d)
    mask = torch.ones(s, s, dtype=torch.bool, device=q.device).tril()
    sc = sc.masked_fill(~mask, float("-inf"))
    p = sc.softmax(-1)
    return (p @ vf).transpose(1, 2)


def swiglu(gate: torch.Tensor, up: torch.Tensor) -> torch.Tensor:
    """silu(gate) * up (reference: model.py:254)."""
    return F.silu(gate) * up


def cross_entropy_sum(logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
    """Sum-reduced CE over fp32 logits with ignore_index=-100 (reference: train.py:101)."""
    return F.cross_entropy(logits.flatten(0, -2).float(), labels.flatten(), reduction="sum")
