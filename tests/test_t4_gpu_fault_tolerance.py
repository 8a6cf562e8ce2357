"""T4/T6 on one GPU: train.py on the native B200 engine through the fake-Slurm harness — SIGUSR1 -> sharded
durable checkpoint -> resubmit -> consolidated reference-layout file -> resume, with a bitwise-identical loss
curve and final state versus an uninterrupted run."""
import os
import signal

import pytest
import torch

from ft_harness import Job
from test_t2_fault_tolerance import _digest

pytestmark = pytest.mark.gpu

GPU_ARGS = ["--tokenizer-name-or-path", "offline", "--vocab-size", "2048", "--dim", "256", "--n-layers", "2",
            "--n-heads", "4", "--n-kv-heads", "2", "--multiple-of", "64", "--sequence-length", "64", "--batch-size", "2",
            "--model-dtype", "bf16", "--device", "cuda", "--synthetic-docs", "64", "--learning-rate", "1e-3",
            "--lr-warmup-steps", "5", "--logging-frequency", "1"]


def _args(tmp, steps, *more):
    return GPU_ARGS + ["--dataset", os.path.join(str(tmp), "data.parquet"), "--training-steps", str(steps)] + list(more)


def test_gpu_usr1_chain_bitwise(tmp_path):
    # enough steps that the job is still running when the signal arrives (a step of this model takes ~2 ms)
    full = Job(tmp_path, 5000, _args(tmp_path, 600, "--save-final"))
    assert full.returncode == 0 and "Training completed" in full.text, full.text[-2000:]
    j1 = Job(tmp_path, 5001, _args(tmp_path, 600), signal_at=8, sig=signal.SIGUSR1)
    assert j1.returncode == 0, j1.text[-2000:]
    assert "[EXIT HANDLER] Job timed out, saving checkpoint." in j1.text
    assert "shards durable in" in j1.text and "consolidated reference-layout file written" in j1.text
    assert j1.sbatch_calls == [f"{tmp_path}/train.sh 5001"]
    # durable (manifest) comes before the resubmit, the consolidated file after it
    assert j1.text.index("shards durable in") < j1.text.index("sbatch requeued") < j1.text.index("consolidated reference")
    j2 = Job(tmp_path, 5002, _args(tmp_path, 600, "--save-final", "--checkpoint_id=5001"))
    assert j2.returncode == 0 and "Training completed" in j2.text, j2.text[-2000:]
    merged = dict(j1.losses); merged.update(j2.losses)
    assert merged == full.losses
    a = torch.load(full.ckpt_path(), weights_only=False)
    b = torch.load(j2.ckpt_path(), weights_only=False)
    assert _digest(a) == _digest(b)


def test_gpu_sigterm_saves_nothing_and_exception_saves(tmp_path):
    j = Job(tmp_path, 6001, _args(tmp_path, 2000), signal_at=3, sig=signal.SIGTERM)
    assert j.returncode == 0 and "[EXIT HANDLER] Job cancelled, terminating." in j.text
    assert not os.path.exists(j.ckpt_path()) and not os.path.isdir(j.ckpt_path().replace(".ckpt", ".shards"))
    k = Job(tmp_path, 6002, _args(tmp_path, 12, "--raise-error", "--error-step", "5", "--no-consolidate"))
    assert "[EXIT HANDLER] Checkpoint saved at step 5" in k.text and k.sbatch_calls == []
    assert os.path.isdir(k.ckpt_path().replace(".ckpt", ".shards")) and not os.path.exists(k.ckpt_path())
    r = Job(tmp_path, 6003, _args(tmp_path, 12, "6002"))   # resume straight from the shard directory
    assert "Resuming training from training_step 5" in r.text and "Training completed" in r.text


def test_gpu_five_job_chain_seq8192_bitwise(tmp_path):
    """BASELINE.json config 5 in miniature: seq 8192, a chain of 5 jobs (4 SIGUSR1 kills, each followed by the
    resubmit that the fake sbatch records), packed dataset — loss curve and final model / optimizer / scheduler /
    dataset state are bitwise those of one uninterrupted run."""
    extra = ["--sequence-length", "8192", "--batch-size", "1", "--dataset-type", "packed", "--vocab-size", "2048"]
    base = [a for a in GPU_ARGS]
    for flag in ("--sequence-length", "--batch-size", "--vocab-size"):
        i = base.index(flag)
        del base[i:i + 2]
    mk = lambda steps, *more: base + extra + ["--dataset", os.path.join(str(tmp_path), "data.parquet"),
                                              "--training-steps", str(steps)] + list(more)
    full = Job(tmp_path, 8000, mk(400, "--save-final"))
    assert "Training completed" in full.text, full.text[-2000:]
    prev, losses = None, {}
    for hop in range(5):
        last = hop == 4
        more = ([str(prev)] if prev else []) + (["--save-final"] if last else [])
        j = Job(tmp_path, 8001 + hop, mk(400, *more), signal_at=None if last else 20 + 80 * hop)
        assert j.returncode == 0, j.text[-2000:]
        if not last:
            assert j.sbatch_calls[-1] == f"{tmp_path}/train.sh {8001 + hop}"
        losses.update(j.losses)
        prev = 8001 + hop
    assert losses == full.losses
    a = torch.load(full.ckpt_path(), weights_only=False)
    b = torch.load(os.path.join(str(tmp_path), "checkpoints", f"checkpoint_{prev}.ckpt"), weights_only=False)
    assert _digest(a) == _digest(b) and a["data_state"] == b["data_state"]
