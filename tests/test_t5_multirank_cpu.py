"""T5 host-side logic on CPU with gloo, world size 2: cross-rank stop agreement, rank-0-only resubmit,
per-rank data position in the checkpoint, and 2 ranks x batch B == 1 rank x batch 2B."""
import os
import signal

import torch

from conftest import TINY_ARGS
from ft_harness import Job


def _args(tmp, steps, *more, batch="2"):
    a = [x for x in TINY_ARGS]
    a[a.index("--batch-size") + 1] = batch
    return a + ["--dataset", os.path.join(str(tmp), "data.parquet"), "--training-steps", str(steps),
                "--logging-frequency", "1"] + list(more)


def test_two_rank_usr1_resume_and_equivalence_with_one_rank(tmp_path):
    full = Job(tmp_path, 7000, _args(tmp_path, 16, "--save-final"), nproc=2)
    assert full.returncode == 0 and "Training completed" in full.text, full.text[-3000:]
    j1 = Job(tmp_path, 7001, _args(tmp_path, 16), nproc=2, signal_at=6, sig=signal.SIGUSR1)
    assert j1.returncode == 0, j1.text[-3000:]
    assert j1.text.count("[EXIT HANDLER] Job timed out, saving checkpoint.") == 1   # rank 0 logs, once
    assert j1.sbatch_calls == [f"{tmp_path}/train.sh 7001"]                          # only rank 0 resubmits
    ck = torch.load(j1.ckpt_path(), weights_only=False)
    stop = ck["training_step"]
    assert isinstance(ck["data_state"], list) and len(ck["data_state"]) == 2
    assert [s["rank"] for s in ck["data_state"]] == [0, 1] and all(s["batches_consumed"] == stop for s in ck["data_state"])
    j2 = Job(tmp_path, 7002, _args(tmp_path, 16, "--save-final", "--checkpoint-id", "7001"), nproc=2)
    assert j2.returncode == 0 and "Training completed" in j2.text, j2.text[-3000:]
    merged = dict(j1.losses); merged.update(j2.losses)
    assert merged == full.losses
    a = torch.load(full.ckpt_path(), weights_only=False)
    b = torch.load(j2.ckpt_path(), weights_only=False)
    for k in a["model"]:
        assert torch.equal(a["model"][k], b["model"][k]), k
    # 2 ranks x batch 2 consume the samples of 1 rank x batch 4, with the loss normalised by the global count
    solo = Job(tmp_path, 7003, _args(tmp_path, 16, batch="4"))
    for step, loss in full.losses.items():
        assert abs(float(loss) - float(solo.losses[step])) < 2e-4, (step, loss, solo.losses[step])


def test_two_rank_sigterm_saves_nothing(tmp_path):
    j = Job(tmp_path, 7101, _args(tmp_path, 500), nproc=2, signal_at=3, sig=signal.SIGTERM)
    assert "[EXIT HANDLER] Job cancelled, terminating." in j.text
    assert not os.path.exists(j.ckpt_path()) and j.sbatch_calls == []


def test_rank_local_failure_stops_every_rank_at_the_same_step(tmp_path):
    """One rank raises on its own: its peers learn about it through the step's single host
    collective, everybody checkpoints the same step, nobody resubmits, and the job resumes."""
    j = Job(tmp_path, 7201, _args(tmp_path, 12), nproc=2, extra_env={"FTL_INJECT_RANK_ERROR": "1:5"})
    assert j.returncode == 0, j.text[-3000:]
    assert "injected rank-local failure on rank 1 at step 5" in j.text
    assert "[EXIT HANDLER] Error during training encountered, saving checkpoint." in j.text
    assert "[EXIT HANDLER] Checkpoint saved at step 5" in j.text and j.sbatch_calls == []
    ck = torch.load(j.ckpt_path(), weights_only=False)
    assert ck["training_step"] == 5 and all(s["batches_consumed"] == 5 for s in ck["data_state"])
    r = Job(tmp_path, 7202, _args(tmp_path, 12, "7201"), nproc=2)
    assert "Resuming training from training_step 5" in r.text and "Training completed" in r.text
    assert set(r.losses) == set(range(5, 12))
