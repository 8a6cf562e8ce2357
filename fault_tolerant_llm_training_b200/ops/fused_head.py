"""LM head + cross-entropy without the [T, V] logits (reference: model.py:379 + train.py:100-103).

The reference materialises bf16 logits, an fp32 copy of them and an fp32 gradient (3.5 GiB at T = 2048, V = 131072).
The default engine path keeps one bf16 [T, V] buffer that the loss kernel turns into d(logits) in place (512 MiB at
seq 2048, 2 GiB at 8192).  This class removes that buffer as well: the vocabulary is processed in chunks of `chunk`
columns through one [T, chunk] buffer that stays in L2.

  forward   per chunk: logits_c = x @ W_c^T (tcgen05 GEMM) -> per-row (max, sum-exp, label logit) of the chunk;
            combine -> row log-sum-exp, loss
  backward  per chunk: recompute logits_c, turn them into d(logits)_c in place, dW_c = d(logits)_c^T @ x (written
            straight into the weight-gradient rows of the chunk, with its gradient-norm partials),
            d(x) += d(logits)_c @ W_c with an fp32 running sum in the GEMM epilogue

It costs one more head GEMM (+2.3 % of a step's FLOPs at the train.sh shape) and saves the buffer plus ~3 GB of HBM
traffic per step, so it is a memory lever, not a speed-up: `FTL_FUSED_HEAD=1` switches it on (default off).
"""
from __future__ import annotations

from typing import Callable, Optional

import torch

from . import native as N


class ChunkedHead:
    def __init__(self, rows: int, dim: int, vocab: int, device: torch.device, chunk: int = 8192):
        if chunk >= vocab or vocab % chunk or chunk % 8:
            chunk = vocab                       # degenerate: one chunk (still correct, no memory saved)
        self.rows, self.dim, self.vocab, self.chunk = rows, dim, vocab, chunk
        self.n_chunks = vocab // chunk
        f32 = lambda *s: torch.empty(*s, dtype=torch.float32, device=device)
        self.buf = torch.empty(rows, chunk, dtype=torch.bfloat16, device=device)
        self.cmax, self.csum = f32(rows, self.n_chunks), f32(rows, self.n_chunks)
        self.picked, self.lse = torch.zeros(rows, dtype=torch.float32, device=device), f32(rows)
        self.dx32 = f32(rows, dim)

    def forward(self, x: torch.Tensor, w: torch.Tensor, labels: torch.Tensor, inv_n: torch.Tensor, row_loss: torch.Tensor,
                loss_out: torch.Tensor) -> torch.Tensor:
        """x: [rows, dim] bf16, w: [vocab, dim] bf16, labels: int64 [rows].  loss_out[0] = sum(CE) * inv_n."""
        C = self.chunk
        for c in range(self.n_chunks):
            N.gemm(x, w[c * C:(c + 1) * C], self.buf)
            N.ce_chunk_stats(self.buf, labels, c * C, self.n_chunks, c, self.cmax, self.csum, self.picked)
        N.ce_combine(self.cmax, self.csum, self.picked, labels, self.n_chunks, self.vocab, inv_n, self.lse, row_loss, loss_out)
        return loss_out

    def backward(self, x: torch.Tensor, w: torch.Tensor, labels: torch.Tensor, inv_n: torch.Tensor, dw: torch.Tensor,
                 dx: torch.Tensor, sumsq: Optional[Callable[[], Optional[torch.Tensor]]] = None) -> None:
        """dw: [vocab, dim] bf16 weight gradient (overwritten); dx: [rows, dim] bf16 = d(loss)/d(x).
        sumsq(): returns the next slice for the weight-gradient GEMM's gradient-norm partials, or None."""
        C, last = self.chunk, self.n_chunks - 1
        for c in range(self.n_chunks):
            wc = w[c * C:(c + 1) * C]
            N.gemm(x, wc, self.buf)                                   # recompute the chunk's logits
            N.ce_chunk_grad(self.buf, labels, c * C, self.vocab, self.lse, inv_n)
            N.gemm(self.buf, x, dw[c * C:(c + 1) * C], a_mn=True, b_mn=True, sumsq=sumsq() if sumsq else None)
            mode = (1 if c > 0 else 0) | (2 if c == last else 0)      # add after the first chunk; bf16 out on the last
            N.gemm_acc32(self.buf, wc, self.dx32, mode, out=dx if c == last else None)
