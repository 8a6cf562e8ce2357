"""Llama-3-style decoder with the reference's module tree and parameter names.

`state_dict()` produces exactly the 291 keys (for the 32-layer config) that
/root/reference/model.py:315-380 produces, in the same order, so a checkpoint
written here loads into the reference `Transformer` with strict=True and the
optimizer's parameter indices line up:

    tok_embeddings.weight
    layers.{i}.attention.{wq,wk,wv,wo}.weight
    layers.{i}.feed_forward.{w1,w2,w3}.weight
    layers.{i}.attention_norm.weight, layers.{i}.ffn_norm.weight
    norm.weight
    output.weight

`forward()` here is the portable PyTorch path (CPU runs, oracles).  On a B200
the same parameters are driven by `models.engine.NativeEngine`, which runs a
hand-scheduled forward/backward over the sm_100a kernels and never calls this
`forward`.
"""
from __future__ import annotations

import math
from typing import Iterator, List, Tuple

import torch
import torch.nn as nn

from ..config import ModelConfig
from ..ops import torch_ops as T


class RMSNorm(nn.Module):
    """Reference: model.py:24-48."""

    def __init__(self, dim: int, eps: float):
        super().__init__()
        self.eps = eps
        self.weight = nn.Parameter(torch.ones(dim))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return T.rmsnorm(x, self.weight, self.eps)


class Attention(nn.Module):
    """Bias-free q/k/v/o projections + RoPE + causal GQA (reference: model.py:141-215)."""

    def __init__(self, cfg: ModelConfig):
        super().__init__()
        self.n_heads, self.n_kv_heads, self.head_dim = cfg.n_heads, cfg.n_kv_heads, cfg.head_dim
        self.wq = nn.Linear(cfg.dim, cfg.n_heads * cfg.head_dim, bias=False)
        self.wk = nn.Linear(cfg.dim, cfg.n_kv_heads * cfg.head_dim, bias=False)
        self.wv = nn.Linear(cfg.dim, cfg.n_kv_heads * cfg.head_dim, bias=False)
        self.wo = nn.Linear(cfg.n_heads * cfg.head_dim, cfg.dim, bias=False)

    def forward(self, x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor) -> torch.Tensor:
        b, s, _ = x.shape
        q = self.wq(x).view(b, s, self.n_heads, self.head_dim)
        k = self.wk(x).view(b, s, self.n_kv_heads, self.head_dim)
        v = self.wv(x).view(b, s, self.n_kv_heads, self.head_dim)
        q, k = T.rope_apply(q, cos, sin), T.rope_apply(k, cos, sin)
        o = T.attention(q, k, v)
        return self.wo(o.reshape(b, s, -1))


class FeedForward(nn.Module):
    """w2(silu(w1 x) * w3 x) (reference: model.py:218-254)."""

    def __init__(self, cfg: ModelConfig):
        super().__init__()
        hidden = cfg.ffn_hidden
        self.w1 = nn.Linear(cfg.dim, hidden, bias=False)
        self.w2 = nn.Linear(hidden, cfg.dim, bias=False)
        self.w3 = nn.Linear(cfg.dim, hidden, bias=False)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.w2(T.swiglu(self.w1(x), self.w3(x)))


class TransformerBlock(nn.Module):
    """Pre-norm residual block (reference: model.py:257-312).  Sub-module
    registration order matches the reference so parameter order does too."""

    def __init__(self, layer_id: int, cfg: ModelConfig):
        super().__init__()
        self.layer_id = layer_id
        self.attention = Attention(cfg)
        self.feed_forward = FeedForward(cfg)
        self.attention_norm = RMSNorm(cfg.dim, cfg.norm_eps)
        self.ffn_norm = RMSNorm(cfg.dim, cfg.norm_eps)

    def forward(self, x, cos, sin):
        h = x + self.attention(self.attention_norm(x), cos, sin)
        return h + self.feed_forward(self.ffn_norm(h))


class Transformer(nn.Module):
    """Reference: model.py:315-380.  The rotary table is a pair of non-persistent
    fp32 buffers (cos, sin) instead of one complex64 buffer; like the reference's
    `freqs_cis` (model.py:342-344) they are not part of the state_dict."""

    def __init__(self, cfg: ModelConfig):
        super().__init__()
        self.cfg = cfg
        self.model_args = cfg  # reference attribute name
        self.vocab_size = cfg.vocab_size
        self.n_layers = cfg.n_layers
        self.tok_embeddings = nn.Embedding(cfg.vocab_size, cfg.dim)
        cos, sin = T.rope_table(cfg.head_dim, cfg.seq_len, cfg.rope_theta)
        self.register_buffer("rope_cos", cos, persistent=False)
        self.register_buffer("rope_sin", sin, persistent=False)
        self.layers = nn.ModuleDict({str(i): TransformerBlock(i, cfg) for i in range(cfg.n_layers)})
        self.norm = RMSNorm(cfg.dim, cfg.norm_eps)
        self.output = nn.Linear(cfg.dim, cfg.vocab_size, bias=False)

    def forward(self, tokens: torch.Tensor) -> torch.Tensor:
        h = self.tok_embeddings(tokens)
        cos, sin = self.rope_cos, self.rope_sin
        for layer in self.layers.values():
            h = layer(h, cos, sin)
        return self.output(self.norm(h))

    # rope tables must stay fp32 when the model is built under a bf16 default dtype
    def _apply(self, fn, recurse=True):
        super()._apply(fn, recurse)
        for name in ("rope_cos", "rope_sin"):
            buf = getattr(self, name)
            if buf.dtype != torch.float32:
                setattr(self, name, buf.float())
        return self


def build_model(cfg: ModelConfig, dtype: torch.dtype, device: torch.device | str = "cpu",
                seed: int = 0) -> Transformer:
    """Create the model with parameters *born* in `dtype` (reference: the
    set_default_dtype context of utils.py:100-110 used at train.py:54) and with
    a fixed seed so every rank initialises identically.

    Initialisation distributions are PyTorch's module defaults, as in the
    reference (it never calls a custom init): kaiming-uniform(a=sqrt(5)) for the
    Linear weights, N(0,1) for the embedding, ones for the norms.  For large
    models the draw happens directly on `device` to avoid the reference's
    35-second CPU init (logs/output_444664.out:5-6).
    """
    device = torch.device(device)
    old = torch.get_default_dtype()
    torch.set_default_dtype(dtype)
    try:
        with torch.device("meta"):
            model = Transformer(cfg)
    finally:
        torch.set_default_dtype(old)
    model.to_empty(device=device)
    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    with torch.no_grad():
        for name, p in model.named_parameters():
            if name.endswith("norm.weight"):
                p.fill_(1.0)
            elif name == "tok_embeddings.weight":
                p.normal_(0.0, 1.0, generator=gen)
            else:
                bound = 1.0 / math.sqrt(p.shape[1])  # kaiming_uniform(a=sqrt(5)) == U(-1/sqrt(fan_in), ..)
                p.uniform_(-bound, bound, generator=gen)
        cos, sin = T.rope_table(cfg.head_dim, cfg.seq_len, cfg.rope_theta, device=device)
        model.rope_cos, model.rope_sin = cos, sin
    return model


def param_layout(model: nn.Module) -> List[Tuple[str, torch.Size, int, int]]:
    """(name, shape, offset, numel) of every parameter in `model.parameters()`
    order, each offset rounded up to 128 elements so that every tensor starts on
    a 256-byte boundary inside a flat bf16 buffer (TMA needs 16-byte alignment;
    256 keeps vector loads and shard boundaries clean)."""
    out, off = [], 0
    for name, p in model.named_parameters():
        n = p.numel()
        out.append((name, p.shape, off, n))
        off += (n + 127) // 128 * 128
    return out


def flat_size(layout) -> int:
    name, shape, off, n = layout[-1]
    return off + (n + 127) // 128 * 128


def iter_layer_param_names(i: int) -> Iterator[str]:
    base = f"layers.{i}."
    for n in ("attention.wq", "attention.wk", "attention.wv", "attention.wo",
              "feed_forward.w1", "feed_forward.w2", "feed_forward.w3",
              "attention_norm", "ffn_norm"):
        yield base + n + ".weight"
