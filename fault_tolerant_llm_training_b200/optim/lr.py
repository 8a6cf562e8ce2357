"""Learning-rate schedule: linear warm-up then constant (reference: utils.py:32-56;
its docstring says "decay" but the code is warm-up-then-constant).

Oracle values for lr=5e-5, warmup=100: steps 0/1/99/100 ->
4.9505e-7 / 9.9010e-7 / 4.9505e-5 / 5e-5.
"""
from __future__ import annotations

import functools

from torch.optim.lr_scheduler import LambdaLR


def warmup_constant_factor(warmup_steps: int, current_step: int) -> float:
    """Multiplicative LR factor at 0-indexed `current_step`."""
    if current_step < warmup_steps:
        return float((current_step + 1) / (warmup_steps + 1))
    return 1.0


def build_lr_scheduler(optimizer, warmup_steps: int) -> LambdaLR:
    """A LambdaLR over a functools.partial, so `state_dict()["lr_lambdas"] == [{}]`
    exactly like the reference's scheduler and either side can load the other's."""
    return LambdaLR(optimizer, functools.partial(warmup_constant_factor, warmup_steps))


def rewind_lr_scheduler(scheduler: LambdaLR, epoch: int) -> None:
    """Put a LambdaLR back to the state it had after `epoch` step() calls (used when steps that the host had
    already counted turn out not to have been applied on the device)."""
    scheduler.last_epoch = epoch
    scheduler._step_count = epoch + 1
    lrs = [base * lmbda(epoch) for base, lmbda in zip(scheduler.base_lrs, scheduler.lr_lambdas)]
    for group, lr in zip(scheduler.optimizer.param_groups, lrs):
        group["lr"] = lr
    scheduler._last_lr = lrs
