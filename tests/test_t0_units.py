"""T0 unit tests: LR schedule oracle values, CLI, hidden-dim rule, error-type mapping,
handle_exit branch table."""
import logging
import os

import pytest
import torch

from fault_tolerant_llm_training_b200.cli import get_args, model_config_from_args
from fault_tolerant_llm_training_b200.config import ModelConfig, swiglu_hidden_dim
from fault_tolerant_llm_training_b200.ft.exit_handler import handle_exit
from fault_tolerant_llm_training_b200.ft.signals import (SignalFlags, TrainingInterrupt, error_type_of,
                                                         STOP_CANCEL, STOP_TIMEOUT)
from fault_tolerant_llm_training_b200.optim.lr import build_lr_scheduler, warmup_constant_factor


def test_lr_schedule_oracle_values():
    p = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.AdamW([p], lr=5e-5)
    sched = build_lr_scheduler(opt, 100)
    seen = {}
    for step in range(102):
        seen[step] = opt.param_groups[0]["lr"]
        opt.step()
        sched.step()
    assert seen[0] == pytest.approx(4.9505e-7, rel=1e-4)
    assert seen[1] == pytest.approx(9.9010e-7, rel=1e-4)
    assert seen[99] == pytest.approx(4.9505e-5, rel=1e-4)
    assert seen[100] == pytest.approx(5e-5, rel=1e-9)
    assert seen[101] == pytest.approx(5e-5, rel=1e-9)
    assert warmup_constant_factor(10, 10) == 1.0
    assert sched.state_dict()["lr_lambdas"] == [{}]  # same as the reference's functools.partial scheduler


def test_lr_scheduler_state_matches_reference(reference_modules):
    p = torch.nn.Parameter(torch.zeros(1))
    ours = build_lr_scheduler(torch.optim.AdamW([p], lr=1e-3), 7)
    theirs = reference_modules["utils"].build_lr_scheduler(torch.optim.AdamW([p], lr=1e-3), 7)
    for s in (ours, theirs):
        for _ in range(11):
            s.optimizer.step()
            s.step()
    assert ours.state_dict() == theirs.state_dict()
    theirs.load_state_dict(ours.state_dict())


def test_hidden_dim_rule():
    assert swiglu_hidden_dim(4096, 1024, 1.3) == 14336
    assert ModelConfig().ffn_hidden == 14336
    assert ModelConfig().num_params() == 8_053_329_920
    assert ModelConfig().qkv_dim == 6144


def test_cli_defaults_match_reference():
    a = get_args([])
    assert (a.sequence_length, a.batch_size, a.learning_rate, a.lr_warmup_steps) == (4096, 1, 1e-5, 10)
    assert (a.training_steps, a.logging_frequency, a.grad_max_norm, a.model_dtype) == (1000, 5, 1, "bf16")
    assert (a.raise_error, a.error_step, a.checkpoint_id, a.compile, a.fused_optimizer) == (False, 100, "", False, False)
    assert a.tokenizer_name_or_path == "unsloth/Mistral-Nemo-Base-2407-bnb-4bit"
    cfg = model_config_from_args(a, 131072)
    assert (cfg.dim, cfg.n_layers, cfg.n_heads, cfg.n_kv_heads, cfg.rope_theta) == (4096, 32, 32, 8, 500000)


def test_cli_aliases_and_positional(monkeypatch):
    a = get_args(["--checkpoint_id=444664", "--raise_error", "--error_step", "600", "--sequence_length", "2048"])
    assert a.checkpoint_id == "444664" and a.raise_error and a.error_step == 600 and a.sequence_length == 2048
    b = get_args(["--checkpoint-id", "7"])
    assert b.checkpoint_id == "7"
    c = get_args(["--training-steps", "5", "444671"])  # train.sh style positional id
    assert c.checkpoint_id == "444671"
    monkeypatch.setenv("WORKDIR", "/some/dir")
    assert get_args([]).checkpoint_path == "/some/dir/checkpoints"


def test_error_type_extraction():
    assert error_type_of(TrainingInterrupt(10)) == 10
    assert error_type_of(TrainingInterrupt(15)) == 15
    assert TrainingInterrupt(10).args == ("Exception", 10)  # the reference's e.args[1] convention
    assert error_type_of(ValueError("x")) == -1
    assert error_type_of(Exception("Simulated exception to test signal handler", -1)) == -1
    assert error_type_of(OSError(28, "No space left on device")) == -1  # reference: "Unknown exit signal"


def test_signal_flags_priority():
    import signal
    f = SignalFlags().install()
    try:
        os.kill(os.getpid(), signal.SIGUSR1)
        assert f.poll() == STOP_TIMEOUT
        os.kill(os.getpid(), signal.SIGTERM)
        assert f.poll() == STOP_CANCEL  # cancel wins over timeout
        os.kill(os.getpid(), signal.SIGUSR1)
        assert f.poll() == STOP_CANCEL
    finally:
        f.uninstall()


class _Recorder:
    def __init__(self):
        self.saved, self.lines = [], []

    def save(self, path, jid, *a, **k):
        self.saved.append((path, jid))
        return os.path.join(path, f"checkpoint_{jid}.ckpt")


@pytest.mark.parametrize("etype,saves,resubmits,needle", [
    (15, 0, 0, "[EXIT HANDLER] Job cancelled, terminating."),
    (10, 1, 1, "[EXIT HANDLER] Job timed out, saving checkpoint."),
    (-1, 1, 0, "[EXIT HANDLER] Error during training encountered, saving checkpoint."),
    (28, 0, 0, "[EXIT HANDLER] Unknown exit signal 28, terminating."),
])
def test_handle_exit_branch_table(tmp_path, monkeypatch, caplog, etype, saves, resubmits, needle):
    rec = _Recorder()
    marker = tmp_path / "resubmitted"
    monkeypatch.setenv("SLURM_JOB_ID", "4242")
    monkeypatch.setenv("WORKDIR", str(tmp_path))
    log = logging.getLogger("t0")
    with caplog.at_level(logging.INFO):
        handle_exit(None, None, None, 37, str(tmp_path), etype, log, save_fn=rec.save,
                    resubmit_cmd=f"echo {{workdir}}/train.sh {{jobid}} > {marker}")
    assert len(rec.saved) == saves
    assert marker.exists() == bool(resubmits)
    if resubmits:
        assert marker.read_text().strip() == f"{tmp_path}/train.sh 4242"
        assert "sbatch requeued, new job will load the last checkpoint" in caplog.text
    if saves:
        assert "[EXIT HANDLER] Checkpoint saved at step 37" in caplog.text
    assert needle in caplog.text


def test_handle_exit_failed_resubmit(tmp_path, monkeypatch, caplog):
    rec = _Recorder()
    monkeypatch.setenv("SLURM_JOB_ID", "9")
    with caplog.at_level(logging.INFO):
        handle_exit(None, None, None, 1, str(tmp_path), 10, logging.getLogger("t0"), save_fn=rec.save,
                    resubmit_cmd="false")
    assert "[EXIT HANDLER] Failed to requeue job 9." in caplog.text


def test_c_header_declares_every_entry_point():
    """include/ftl_b200.h is the documented C ABI: every exported function of csrc/ must be declared."""
    import glob
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    header = open(os.path.join(root, "include", "ftl_b200.h")).read()
    declared = set(re.findall(r"\b(ftl_[a-z0-9_]+)\s*\(", header))
    defined = set()
    for path in glob.glob(os.path.join(root, "csrc", "*.cu")):
        defined |= set(re.findall(r"^FTL_API\s+[\w\s\*]+?\b(ftl_[a-z0-9_]+)\s*\(", open(path).read(), flags=re.M))
    defined.discard("ftl_attn_dbg_read")            # probe build only (FTL_ATTN_DEBUG)
    assert len(defined) > 40 and not (defined - declared), sorted(defined - declared)


def test_loader_picks_the_newest_complete_shard_directory(tmp_path):
    """Crash-safe replacement of a sharded checkpoint: whichever of `.shards`, `.shards.new`, `.shards.old`
    is complete (has a manifest) and has the largest training_step wins; an incomplete directory is ignored."""
    import json

    from fault_tolerant_llm_training_b200.ckpt.layout import load_checkpoint
    from fault_tolerant_llm_training_b200.ckpt.sharded import MANIFEST, best_shard_dir, shard_dir

    base = shard_dir(str(tmp_path), "77")

    def write(d, step, complete=True):
        os.makedirs(d, exist_ok=True)
        open(os.path.join(d, "rank0.bin"), "wb").close()
        if complete:
            with open(os.path.join(d, MANIFEST), "w") as f:
                json.dump({"training_step": step, "lr_scheduler": {}, "world_size": 1}, f)

    assert best_shard_dir(str(tmp_path), "77") is None
    write(base, 10)
    assert best_shard_dir(str(tmp_path), "77") == base
    write(base + ".new", 20, complete=False)          # killed while the new shards were being written
    assert best_shard_dir(str(tmp_path), "77") == base
    write(base + ".new", 20)                           # manifest durable, killed before the swap
    assert best_shard_dir(str(tmp_path), "77") == base + ".new"
    ck = load_checkpoint(str(tmp_path), "77")          # no consolidated file: the loader resumes from the newest shards
    assert ck["training_step"] == 20 and ck["_sharded_dir"] == base + ".new"
    os.rename(base, base + ".old")                     # killed inside the swap: only .old and .new exist
    assert best_shard_dir(str(tmp_path), "77") == base + ".new"
