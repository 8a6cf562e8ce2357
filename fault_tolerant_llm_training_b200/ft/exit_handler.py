"""Exit handler: the reference's three interruption paths (utils.py:65-90, README.md:33):

    type 15  (SIGTERM / scancel)  -> log only, nothing saved
    type 10  (SIGUSR1 / timeout)  -> save checkpoint, then resubmit
    type -1  (Python exception)   -> save checkpoint, no resubmit
    other                         -> "Unknown exit signal", nothing saved

The log strings are the reference's, byte for byte (users and tests grep them).
The save itself is delegated to `ckpt.save_checkpoint`, which writes the
reference's torch.save layout atomically and, on multi-rank runs, per-rank
shards first.  Only rank 0 resubmits.
"""
from __future__ import annotations

import os
import subprocess
import time
from typing import Callable, Optional

from ..config import job_id, workdir

DEFAULT_RESUBMIT = "sbatch {workdir}/train.sh {jobid}"


def resubmit(jobid: str, template: str = "", logger=None) -> int:
    """Run the resubmit command (reference: os.system(f"sbatch {WORKDIR}/train.sh {JOBID}"),
    utils.py:84).  Returns the exit status."""
    cmd = (template or os.environ.get("FTL_RESUBMIT_CMD") or DEFAULT_RESUBMIT).format(
        workdir=workdir(), jobid=jobid)
    try:
        return subprocess.call(cmd, shell=True)
    except OSError as e:
        if logger is not None:
            logger.info(f"[EXIT HANDLER] resubmit command failed to start: {e}")
        return 127


def handle_exit(model, optimizer, lr_scheduler, training_step, checkpoint_path, type, logger,
                *, data_state: Optional[dict] = None, rank: int = 0, world_size: int = 1,
                save_fn: Optional[Callable[..., str]] = None, resubmit_cmd: str = "",
                resume_id: str = "", made_progress: bool = True, flags=None,
                extra: Optional[dict] = None, finalize_fn: Optional[Callable[..., object]] = None) -> Optional[str]:
    """Same positional signature as the reference's handle_exit (utils.py:65).

    Keyword-only extensions: dataset position, rank info, a pluggable save
    function (sharded engines pass theirs), the resubmit command template, and
    `made_progress`/`resume_id` for a timeout that arrives before the first step
    of this job (then the job is resubmitted against the checkpoint it started
    from and nothing is written: SURVEY Q12).  Returns the checkpoint path or None.
    """
    info = logger.info if rank == 0 else (lambda *_a, **_k: None)
    jid = job_id()
    saved_path = None
    if type == 15:
        info("[EXIT HANDLER] Job cancelled, terminating.")
    elif type in (-1, 10):
        if type == 10:
            info("[EXIT HANDLER] Job timed out, saving checkpoint.")
        else:
            info("[EXIT HANDLER] Error during training encountered, saving checkpoint.")
        next_id = jid
        if made_progress or not resume_id or type == -1:
            if flags is not None:
                flags.saving = True
            t0 = time.perf_counter()
            if save_fn is None:
                from ..ckpt.layout import save_checkpoint as save_fn  # noqa: PLW0642
            saved_path = save_fn(checkpoint_path, jid, model, optimizer, lr_scheduler, training_step,
                                 data_state=data_state, rank=rank, world_size=world_size, extra=extra,
                                 logger=logger)
            if flags is not None:
                flags.saving = False
            info(f"[EXIT HANDLER] Checkpoint saved at step {training_step}")
            info(f"[EXIT HANDLER] checkpoint wall time {time.perf_counter() - t0:.2f} s -> {saved_path}")
        else:
            next_id = resume_id
            info(f"[EXIT HANDLER] No step completed in this job; keeping checkpoint {resume_id}")
        resubmitted = False
        if type == 10 and rank == 0:
            ret = resubmit(next_id, resubmit_cmd, logger)
            resubmitted = ret == 0
            if ret != 0:
                info(f"[EXIT HANDLER] Failed to requeue job {jid}.")
            else:
                info("[EXIT HANDLER] sbatch requeued, new job will load the last checkpoint")
        if saved_path is not None and finalize_fn is not None:
            # work that may run after the job has been resubmitted (sharded engines consolidate the shards into
            # the reference-layout file here); the checkpoint was already durable when save_fn returned
            import inspect
            if "resubmitted" in inspect.signature(finalize_fn).parameters:
                finalize_fn(logger=logger, resubmitted=resubmitted)
            else:
                finalize_fn(logger=logger)
    else:
        info(f"[EXIT HANDLER] Unknown exit signal {type}, terminating.")
    return saved_path
