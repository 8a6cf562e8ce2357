"""Training driver — the counterpart of /root/reference/train.py:12-129.

Same shape (setup -> optional resume -> loop -> every interruption routed to
`handle_exit` -> exit code 0), same log lines.  Differences, each from a numbered
quirk in SURVEY.md §9:

  * signal handlers are installed before setup and only set a flag; the loop acts
    on it at the step boundary after all ranks agree, so a checkpoint never holds
    a half-applied optimizer step (Q2, Q12);
  * the injected error fires at the top of step `--error-step`, so "saved at step
    N" means exactly N updates were applied (Q3);
  * any non-signal exception is error type -1 (Q4);
  * checkpoints are loaded strictly and saved atomically with the dataset
    position (Q5, Q11); resume restores the data position in O(1);
  * a seed makes all ranks (and resumed jobs) initialise identically (Q7).
"""
from __future__ import annotations

import os
import sys
import time

import torch

from .cli import model_config_from_args
from .ckpt.layout import load_checkpoint
from .config import PRECISION_STR_TO_DTYPE
from .data.datasets import BatchStream, CollatorForCLM, IterableParquetDataset, ParquetDataset
from .data.synthetic import write_synthetic_parquet
from .data.tokenizer import load_tokenizer
from .ft.exit_handler import handle_exit
from .ft.signals import (STOP_ERROR, STOP_NONE, SignalFlags, TrainingInterrupt, error_type_of,
                         stop_code_to_error_type)
from .log import logger
from .parallel.env import init_dist
from .trainer import make_trainer


def _build_stream(args, tokenizer, env) -> BatchStream:
    if not os.path.exists(args.dataset) and args.synthetic_docs > 0:
        if env.is_main:
            write_synthetic_parquet(args.dataset, num_docs=args.synthetic_docs,
                                    max_len=max(128, 2 * args.sequence_length),
                                    vocab_size=tokenizer.vocab_size)
        env.barrier()
    if args.dataset_type == "packed":
        ds = IterableParquetDataset(args.dataset, tokenizer, args.sequence_length,
                                    bos_token_id=getattr(tokenizer, "bos_token_id", 1) or 1,
                                    offset=env.rank, stride=env.world_size)
        return BatchStream(ds, args.batch_size, rank=env.rank, world_size=env.world_size,
                           pin_memory=env.device.type == "cuda")
    ds = ParquetDataset(args.dataset, tokenizer, args.sequence_length,
                        args.batch_size * args.training_steps * env.world_size)
    collator = CollatorForCLM(args.sequence_length, tokenizer.pad_token_id)
    return BatchStream(ds, args.batch_size, collator, rank=env.rank, world_size=env.world_size,
                       pin_memory=env.device.type == "cuda")


def train(args) -> None:
    flags = SignalFlags().install()  # before setup
    env = init_dist(args.device)
    logger.info(f"Experiment args: {args}")
    model_dtype = PRECISION_STR_TO_DTYPE[args.model_dtype]

    checkpoint = None
    if args.checkpoint_id:
        logger.info(f"Loading checkpoint from {args.checkpoint_path}")
        t0 = time.perf_counter()
        checkpoint = load_checkpoint(args.checkpoint_path, args.checkpoint_id)
        logger.info(f"Checkpoint read in {time.perf_counter() - t0:.2f} s")

    logger.info("Setting up DataLoaders...")
    tokenizer = load_tokenizer(args.tokenizer_name_or_path, args.vocab_size, logger)
    stream = _build_stream(args, tokenizer, env)
    if checkpoint is not None:
        states = checkpoint.get("data_state")
        if states is not None:
            stream.restore(states)   # validates world size / batch size against the checkpoint (or converts, padded mode)
        else:  # file written by the reference: position = training_step batches (train.py:36-39)
            stream.fast_forward(int(checkpoint["training_step"]))

    logger.info("Setting up Model...")
    cfg = model_config_from_args(args, tokenizer.vocab_size)
    trainer = make_trainer(cfg, env, dtype=model_dtype, lr=args.learning_rate,
                           warmup_steps=args.lr_warmup_steps, grad_max_norm=args.grad_max_norm,
                           clip_mode=args.clip_mode, seed=args.seed, parallel=args.parallel,
                           fused_optimizer=args.fused_optimizer,
                           keep_shards=args.sharded_checkpoint,
                           consolidate=not args.no_consolidate, compile_model=args.compile)
    if args.compile:
        logger.info("Using `torch.compile`" if env.device.type != "cuda" else
                    "`--compile` ignored: the B200 engine runs hand-written kernels, not a traced graph")
    if checkpoint is not None:
        trainer.load_state(checkpoint)
        if hasattr(trainer, "checkpointer"):
            # its shards may still be read by nobody but us from now on: removed after OUR first durable checkpoint
            trainer.checkpointer.predecessor = (args.checkpoint_path, args.checkpoint_id)
        logger.info("Model loaded from checkpoint")
        logger.info("Optimizer loaded from checkpoint")
        logger.info("LR Scheduler loaded from checkpoint")
        training_step = int(checkpoint["training_step"])
        logger.info(f"Resuming training from training_step {training_step}")
    else:
        training_step = 0
        logger.info("Starting training!")
    del checkpoint
    first_step = training_step
    steps_base = getattr(trainer.optimizer, "steps_done", 0)
    batches_base = stream.batches_consumed

    tokens_per_step = args.batch_size * args.sequence_length * env.world_size
    t_window, window_steps = time.perf_counter(), 0
    # Device-side step timer: a CUDA event is recorded on the compute stream at every logging step; the
    # time between the two most recent COMPLETED events is reported one window late, so the host never waits for it.
    on_gpu = env.device.type == "cuda"
    dev_marks = []   # [(event, steps in the window that ended there)]
    nvtx = torch.cuda.nvtx if on_gpu else None
    exit_kw = dict(rank=env.rank, world_size=env.world_size, resubmit_cmd=args.resubmit_cmd,
                   resume_id=args.checkpoint_id, flags=flags)
    # FTL_INJECT_RANK_ERROR="rank:step" raises on one rank only (test hook for the peer-failure path)
    inject = os.environ.get("FTL_INJECT_RANK_ERROR", "")
    inject_rank_error = tuple(int(x) for x in inject.split(":")) if inject else None
    agreed = False
    try:
        while training_step < args.training_steps:
            # ---- step boundary: the only place an interruption is acted on ----
            if args.raise_error and training_step == args.error_step:
                # fired *before* the step, so the saved state has exactly `error_step` updates;
                # a job resumed from that very checkpoint does not re-fire on its first step
                if not (args.checkpoint_id and training_step == first_step):
                    raise Exception("Simulated exception to test signal handler", -1)
            if inject_rank_error == (env.rank, training_step):
                raise RuntimeError(f"injected rank-local failure on rank {env.rank} at step {training_step}")
            if args.checkpoint_every and training_step > first_step and training_step % args.checkpoint_every == 0:
                # the save is a sequence of host collectives: agree first, so that a rank that failed on its own
                # meets a step_sync here and not a gather it can never pair with
                code, _ = env.step_sync(flags.poll(), 0)
                if code != STOP_NONE:
                    agreed = True
                    raise TrainingInterrupt(stop_code_to_error_type(code))
                _save_now(trainer, args, training_step, stream, env)
                logger.info(f"Periodic checkpoint saved at step {training_step}")

            input_ids, labels = next(stream)
            # the step's single host collective: stop-flag agreement + global token count
            code, n_items = env.step_sync(flags.poll(), int(labels.ne(-100).sum()))
            if code != STOP_NONE:
                agreed = True
                raise TrainingInterrupt(stop_code_to_error_type(code))
            if nvtx is not None:
                nvtx.range_push(f"step {training_step}")
            loss = trainer.step(input_ids, labels, n_items=n_items)
            if nvtx is not None:
                nvtx.range_pop()

            window_steps += 1
            if training_step == 1 or training_step % args.logging_frequency == 0:
                if on_gpu:
                    ev = torch.cuda.Event(enable_timing=True)
                    ev.record()
                    dev_marks.append((ev, window_steps))
                loss_val = trainer.loss_value(loss)  # device -> host sync, as loss.item() in train.py:116
                now = time.perf_counter()
                ms = (now - t_window) * 1e3 / max(window_steps, 1)
                dev = ""
                if len(dev_marks) >= 3 and dev_marks[-2][0].query():   # previous window: both of its events are done
                    dms = dev_marks[-3][0].elapsed_time(dev_marks[-2][0]) / max(dev_marks[-2][1], 1)
                    dev = f" | device {dms:.1f} ms/step (previous window, CUDA events)"
                    del dev_marks[:-2]
                logger.info(f"Training step: {training_step} | Loss: {loss_val:.2f}")
                logger.info(f"  step {training_step}: loss {loss_val:.6f} | {ms:.1f} ms/step | "
                            f"{tokens_per_step / max(ms, 1e-9) * 1e3:.0f} tokens/s | "
                            f"lr {trainer.optimizer.param_groups[0]['lr']:.4e} | "
                            f"grad_norm {float(trainer.last_grad_norm):.4f}{dev}")
                t_window, window_steps = time.perf_counter(), 0
            training_step += 1
        # One more agreement after the last step: a rank that fails locally during its final step (or while its peers
        # are already past the loop) pairs its step_sync(STOP_ERROR) with this one instead of hanging on a collective
        # that nobody else enters.
        # (Signals of this rank are not offered here: a timeout signal that arrives after the last step changes nothing.)
        code, _ = env.step_sync(STOP_NONE, 0)
        if code != STOP_NONE:
            agreed = True
            raise TrainingInterrupt(stop_code_to_error_type(code))
        logger.info("Training completed")
        if args.save_final:
            _save_now(trainer, args, training_step, stream, env)
            logger.info(f"Final checkpoint saved at step {training_step}")
        elif hasattr(trainer, "checkpointer"):
            trainer.checkpointer.drop_predecessor(logger)   # nothing of ours replaces them: remove the shards we resumed from
    except Exception as e:  # noqa: BLE001 — every interruption goes through the same handler (train.py:121-129)
        error_type = error_type_of(e)
        if error_type == -1 and not (len(e.args) >= 2 and e.args[1] == -1):
            logger.exception("Exception in the training loop")
        if env.distributed and not agreed:
            # this rank left the loop on its own (local exception, injected error, PeerAbort): pair with the peers'
            # next step_sync so that they stop at the same boundary instead of waiting for this rank forever
            if hasattr(trainer, "abort_peers") and error_type == -1:
                trainer.abort_peers()      # peers' kernels parked in a device barrier fall through (status 2)
            env.step_sync(STOP_ERROR, 0)
        training_step = _settle(trainer, stream, e, training_step, first_step, steps_base, batches_base, env)
        handle_exit(trainer.model, trainer.optimizer, trainer.lr_scheduler, training_step,
                    args.checkpoint_path, error_type, logger,
                    data_state=_data_state(stream, env) if error_type in (-1, 10) else None,
                    save_fn=trainer.save_fn(), finalize_fn=trainer.finalize_fn(),
                    made_progress=training_step > first_step, **exit_kw)
    finally:
        stream.close()
        trainer.close()
        flags.uninstall()
        env.shutdown()
    sys.exit(0)  # exit code 0 on every path, as the reference (train.py:119, 129)


def _settle(trainer, stream, exc, training_step, first_step, steps_base, batches_base, env) -> int:
    """Bring the host-side counters in line with what the device actually applied before a checkpoint is taken.

    The host enqueues steps ahead of the GPU.  A step whose gradient norm was non-finite, or that was cut short
    by a failing peer (abort word / barrier timeout), leaves parameters and moments untouched on the device,
    but the host already counted it.  The update kernels count what they applied; everything else is rewound to
    that count, so "saved at step N" again means exactly N updates."""
    from .ft.signals import PeerAbort, TrainingInterrupt
    from .optim.lr import rewind_lr_scheduler

    if not hasattr(trainer, "committed_updates"):
        stream.rewind(batches_base + (training_step - first_step))   # un-fetch the batch of the step that never ran
        return training_step
    done = first_step + trainer.committed_updates()
    done = env.max_int(-env.max_int(-done)) if env.distributed else done   # min over ranks (they agree; belt and braces)
    if done != training_step or trainer.optimizer.steps_done != steps_base + (done - first_step):
        logger.info(f"[EXIT HANDLER] device applied {done - first_step} updates in this job; host had counted "
                    f"{training_step - first_step}: rewinding to step {done}")
        trainer.optimizer.steps_done = steps_base + (done - first_step)
        rewind_lr_scheduler(trainer.lr_scheduler, done)
    stream.rewind(batches_base + (done - first_step))   # a batch fetched for a step that never committed is re-served
    return done


def _save_now(trainer, args, training_step, stream, env) -> None:
    """Checkpoint outside the exit handler (periodic / final): same writers, same layout."""
    from .ckpt.layout import save_checkpoint
    from .config import job_id

    if hasattr(trainer, "drain"):
        trainer.drain()
    (trainer.save_fn() or save_checkpoint)(
        args.checkpoint_path, job_id(), trainer.model, trainer.optimizer, trainer.lr_scheduler, training_step,
        data_state=_data_state(stream, env), rank=env.rank, world_size=env.world_size, logger=logger)
    fin = trainer.finalize_fn()
    if fin is not None:
        fin(logger=logger)


def _data_state(stream: BatchStream, env):
    st = stream.state_dict()
    if env.distributed:
        return env.gather_objects(st)
    return st
