"""Gradient-norm and clipping for the portable (PyTorch) path.

The reference computes the norm, raises on a non-finite value, and then hands an
already-exhausted generator to clip_grads_with_norm_, so no gradient is ever
scaled (utils.py:58-63 called at train.py:107).  `mode="clip"` does
the clipping the reference intended; `mode="reference"` reproduces what it does.
"""
from __future__ import annotations

from typing import Iterable

import torch


@torch.no_grad()
def clip_grad_norm_(parameters: Iterable[torch.nn.Parameter], grad_max_norm: float,
                    mode: str = "clip") -> torch.Tensor:
    params = [p for p in parameters if p.grad is not None]
    grads = [p.grad for p in params]
    total_norm = torch.nn.utils.get_total_norm(grads, error_if_nonfinite=True)
    if mode == "clip":
        torch.nn.utils.clip_grads_with_norm_(params, grad_max_norm, total_norm)
    return total_norm
