"""Model / run configuration.

The reference hard-codes its 8.05 B-parameter architecture inside the driver
(/root/reference/train.py:43-53) and only exposes a dataclass with different
defaults (/root/reference/model.py:9-21).  Here the *defaults* of ModelConfig are
the train.py values, and every field is reachable from the CLI so the tiny CPU
configuration (2 layers, dim 256) of BASELINE.json config 1 needs no code edits.
"""
from __future__ import annotations

import os
from dataclasses import asdict, dataclass, field
from typing import Optional

import torch

# Same spelling table as /root/reference/utils.py:14-19.
PRECISION_STR_TO_DTYPE = {
    "fp16": torch.float16,
    "bf16": torch.bfloat16,
    "fp32": torch.float32,
    "fp64": torch.float64,
}


def swiglu_hidden_dim(dim: int, multiple_of: int, ffn_dim_multiplier: Optional[float]) -> int:
    """SwiGLU hidden width rule (reference: model.py:243-247 with hidden_dim=4*dim,
    model.py:279).  4096 / 1.3 / 1024 -> 14336."""
    hidden = int(2 * (4 * dim) / 3)
    if ffn_dim_multiplier is not None:
        hidden = int(ffn_dim_multiplier * hidden)
    return multiple_of * ((hidden + multiple_of - 1) // multiple_of)


@dataclass
class ModelConfig:
    """Decoder shape.  Defaults reproduce /root/reference/train.py:43-53."""

    dim: int = 4096
    n_layers: int = 32
    n_heads: int = 32
    n_kv_heads: Optional[int] = 8
    multiple_of: int = 1024
    ffn_dim_multiplier: Optional[float] = 1.3
    norm_eps: float = 1e-5
    rope_theta: float = 500000.0
    seq_len: int = 2048
    vocab_size: int = 131072

    def __post_init__(self) -> None:
        if self.n_kv_heads is None:
            self.n_kv_heads = self.n_heads
        if self.dim % self.n_heads:
            raise ValueError(f"dim {self.dim} not divisible by n_heads {self.n_heads}")
        if self.n_heads % self.n_kv_heads:
            raise ValueError("n_heads must be a multiple of n_kv_heads")
        if (self.dim // self.n_heads) % 2:
            raise ValueError("head_dim must be even for rotary embeddings")

    @property
    def head_dim(self) -> int:
        return self.dim // self.n_heads

    @property
    def n_rep(self) -> int:
        return self.n_heads // self.n_kv_heads

    @property
    def ffn_hidden(self) -> int:
        return swiglu_hidden_dim(self.dim, self.multiple_of, self.ffn_dim_multiplier)

    @property
    def qkv_dim(self) -> int:
        return (self.n_heads + 2 * self.n_kv_heads) * self.head_dim

    def num_params(self) -> int:
        d, f, v = self.dim, self.ffn_hidden, self.vocab_size
        per_layer = d * self.qkv_dim + d * d + 3 * d * f + 2 * d
        return 2 * v * d + self.n_layers * per_layer + d

    def flops_per_token_fwd(self, seq_len: Optional[int] = None) -> float:
        """Forward FLOPs per token (GEMMs + causal attention), SURVEY.md §6 counting."""
        s = seq_len or self.seq_len
        d, f, v = self.dim, self.ffn_hidden, self.vocab_size
        gemm = 2 * (self.n_layers * (d * self.qkv_dim + d * d + 3 * d * f) + v * d)
        attn = self.n_layers * 2 * 2 * s * d / 2  # QK^T and PV, causal half
        return gemm + attn

    def to_dict(self) -> dict:
        return asdict(self)


def job_id() -> str:
    """Slurm job id, read at *call* time (the reference reads it at import,
    utils.py:12, and writes checkpoint_None.ckpt without Slurm — SURVEY Q10)."""
    jid = os.environ.get("SLURM_JOB_ID")
    if jid:
        return jid
    jid = os.environ.get("FTL_JOB_ID")
    if jid:
        return jid
    return f"local{os.getpid()}"


def workdir() -> str:
    """WORKDIR, read at call time (reference: utils.py:11)."""
    return os.getenv("WORKDIR", "")
