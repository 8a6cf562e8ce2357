"""Command line.

Every flag of the reference parser (/root/reference/utils.py:112-203) is kept with
the same name, type and default.  Additions: an underscore alias
for every hyphenated flag (the reference README spells `--checkpoint_id`,
`--raise_error`, `--error_step`: README.md:53-55), a positional job id as in
`train.sh <JOBID>` (train.sh:24-27), model-shape flags whose defaults are the
values hard-coded at train.py:43-53, parallel mode, seed and offline-data options.
"""
from __future__ import annotations

import argparse
from typing import List, Optional

from .config import ModelConfig, PRECISION_STR_TO_DTYPE, workdir

DEFAULT_DATASET = "/capstor/store/cscs/ethz/large-sc/datasets/train_data.parquet"
DEFAULT_TOKENIZER = "unsloth/Mistral-Nemo-Base-2407-bnb-4bit"


def _flag(parser: argparse.ArgumentParser, name: str, **kw) -> None:
    """Register `--some-flag` together with its `--some_flag` spelling."""
    names = [name]
    alt = "--" + name[2:].replace("-", "_")
    if alt != name:
        names.append(alt)
    parser.add_argument(*names, dest=name[2:].replace("-", "_"), **kw)


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(
        description="B200-native fault-tolerant LLM pretraining (reference-compatible CLI)"
    )
    # ---- the 15 reference flags (utils.py:114-201) + checkpoint-id ----
    _flag(p, "--dataset", type=str, default=DEFAULT_DATASET,
          help="Parquet file with a 'text' column of documents")
    _flag(p, "--checkpoint-path", type=str, default=None,
          help="Directory for checkpoints (default: $WORKDIR/checkpoints)")
    _flag(p, "--checkpoint-id", type=str, default="",
          help="Job id of the checkpoint to resume from")
    _flag(p, "--tokenizer-name-or-path", type=str, default=DEFAULT_TOKENIZER,
          help="HF tokenizer id/path, or 'offline' for the built-in hashing tokenizer")
    _flag(p, "--sequence-length", type=int, default=4096)
    _flag(p, "--batch-size", type=int, default=1, help="Per-rank batch size")
    _flag(p, "--fused-optimizer", action="store_true",
          help="Accepted for compatibility; the CUDA path is always fused")
    _flag(p, "--learning-rate", type=float, default=1e-5)
    _flag(p, "--lr-warmup-steps", type=int, default=10)
    _flag(p, "--training-steps", type=int, default=1000)
    _flag(p, "--logging-frequency", type=int, default=5,
          help="Log every `--logging-frequency` steps")
    _flag(p, "--grad-max-norm", type=float, default=1)
    _flag(p, "--model-dtype", type=str, default="bf16", choices=sorted(PRECISION_STR_TO_DTYPE),
          help="dtype of parameters, gradients and optimizer states")
    _flag(p, "--compile", action="store_true",
          help="Accepted for compatibility; no tracing compiler is used on the B200 path")
    _flag(p, "--raise-error", action="store_true",
          help="Raise a simulated error at --error-step (fault injection)")
    _flag(p, "--error-step", type=int, default=100)
    p.add_argument("positional_checkpoint_id", nargs="?", default=None,
                   help="Job id to resume from (same as --checkpoint-id; train.sh style)")

    # ---- model shape: defaults are train.py:43-53 ----
    d = ModelConfig()
    _flag(p, "--dim", type=int, default=d.dim)
    _flag(p, "--n-layers", type=int, default=d.n_layers)
    _flag(p, "--n-heads", type=int, default=d.n_heads)
    _flag(p, "--n-kv-heads", type=int, default=d.n_kv_heads)
    _flag(p, "--ffn-dim-multiplier", type=float, default=d.ffn_dim_multiplier)
    _flag(p, "--multiple-of", type=int, default=d.multiple_of)
    _flag(p, "--rope-theta", type=float, default=d.rope_theta)
    _flag(p, "--norm-eps", type=float, default=d.norm_eps)
    _flag(p, "--vocab-size", type=int, default=0,
          help="0 = take it from the tokenizer (reference behaviour, train.py:51)")

    # ---- new: execution ----
    _flag(p, "--parallel", type=str, default="auto", choices=["auto", "none", "dp", "fsdp"],
          help="auto = none at world size 1, dp otherwise")
    _flag(p, "--device", type=str, default="auto", help="auto | cpu | cuda")
    _flag(p, "--seed", type=int, default=0, help="Init seed; identical on every rank (SURVEY Q7)")
    _flag(p, "--clip-mode", type=str, default="clip", choices=["clip", "reference"],
          help="clip = real gradient clipping; reference = norm + finite check only, "
               "which is what the reference actually executes (SURVEY Q1)")
    _flag(p, "--dataset-type", type=str, default="padded", choices=["padded", "packed"],
          help="padded = ParquetDataset+CollatorForCLM (what train.py uses); "
               "packed = IterableParquetDataset")
    _flag(p, "--synthetic-docs", type=int, default=0,
          help="If >0 and --dataset does not exist, write a synthetic parquet with this many docs")
    _flag(p, "--resubmit-cmd", type=str, default="",
          help="Command template for the SIGUSR1 resubmit; default 'sbatch {workdir}/train.sh {jobid}'")
    _flag(p, "--sharded-checkpoint", action="store_true",
          help="Also keep the per-rank shard files after the consolidated file is written")
    _flag(p, "--no-consolidate", action="store_true",
          help="Write per-rank shards only (fastest durable save); resume reads the shards")
    _flag(p, "--checkpoint-every", type=int, default=0,
          help="Periodic checkpoint interval in steps (0 = only on interruption, as the reference)")
    _flag(p, "--save-final", action="store_true",
          help="Also write a checkpoint when training completes (the reference writes none)")
    return p


def get_args(argv: Optional[List[str]] = None) -> argparse.Namespace:
    """Parse the command line (reference: utils.py:112 `get_args`)."""
    args = build_parser().parse_args(argv)
    if args.positional_checkpoint_id and not args.checkpoint_id:
        args.checkpoint_id = args.positional_checkpoint_id
    del args.positional_checkpoint_id
    if args.checkpoint_path is None:
        # reference default f"{WORKDIR}/checkpoints" (utils.py:123), resolved at call time
        args.checkpoint_path = f"{workdir()}/checkpoints"
    return args


def model_config_from_args(args: argparse.Namespace, vocab_size: int) -> ModelConfig:
    return ModelConfig(
        dim=args.dim,
        n_layers=args.n_layers,
        n_heads=args.n_heads,
        n_kv_heads=args.n_kv_heads,
        multiple_of=args.multiple_of,
        ffn_dim_multiplier=args.ffn_dim_multiplier,
        norm_eps=args.norm_eps,
        rope_theta=args.rope_theta,
        seq_len=args.sequence_length,
        vocab_size=args.vocab_size or vocab_size,
    )
