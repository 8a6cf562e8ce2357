"""T2 (BASELINE.json config 1): the three interruption paths and resume, end to end through train.py on CPU,
with a fake sbatch.  Tiny model, synthetic parquet, fp32 so that loss sequences compare exactly."""
import hashlib
import os
import signal

import torch

from conftest import TINY_ARGS
from ft_harness import Job


def _args(tmp, steps, *more):
    return TINY_ARGS + ["--dataset", os.path.join(str(tmp), "data.parquet"), "--training-steps", str(steps),
                        "--logging-frequency", "1"] + list(more)


def _digest(ckpt):
    h = hashlib.sha256()
    for k, v in ckpt["model"].items():
        h.update(k.encode()); h.update(v.reshape(-1).contiguous().view(torch.uint8).numpy().tobytes())
    for i in sorted(ckpt["optimizer"]["state"]):
        for name in ("step", "exp_avg", "exp_avg_sq"):
            h.update(ckpt["optimizer"]["state"][i][name].reshape(-1).contiguous().view(torch.uint8).numpy().tobytes())
    h.update(repr(ckpt["lr_scheduler"]["last_epoch"]).encode())
    h.update(repr(ckpt["training_step"]).encode())
    return h.hexdigest()


def test_sigusr1_checkpoints_resubmits_and_resume_matches_uninterrupted(tmp_path):
    full = Job(tmp_path, 1000, _args(tmp_path, 20, "--save-final"))
    assert full.returncode == 0 and "Training completed" in full.text
    assert len(full.losses) == 20

    j1 = Job(tmp_path, 1001, _args(tmp_path, 20), signal_at=10, sig=signal.SIGUSR1)
    assert j1.returncode == 0
    assert "[EXIT HANDLER] Job timed out, saving checkpoint." in j1.text
    assert "sbatch requeued, new job will load the last checkpoint" in j1.text
    assert j1.sbatch_calls == [f"{tmp_path}/train.sh 1001"]  # reference: sbatch {WORKDIR}/train.sh {JOBID}
    ck = torch.load(j1.ckpt_path(), weights_only=False)
    stop = ck["training_step"]
    assert 10 <= stop < 20 and ck["data_state"]["batches_consumed"] == stop
    assert f"Checkpoint saved at step {stop}" in j1.text

    j2 = Job(tmp_path, 1002, _args(tmp_path, 20, "--save-final", "--checkpoint_id=1001"))
    assert j2.returncode == 0 and f"Resuming training from training_step {stop}" in j2.text
    merged = dict(j1.losses); merged.update(j2.losses)
    assert set(j2.losses) == set(range(stop, 20))
    assert merged == full.losses  # identical loss curve across kill/resume
    a = torch.load(full.ckpt_path(), weights_only=False)
    b = torch.load(j2.ckpt_path(), weights_only=False)
    assert _digest(a) == _digest(b)  # bitwise-identical model / optimizer / scheduler state


def test_sigterm_logs_and_saves_nothing(tmp_path):
    j = Job(tmp_path, 2001, _args(tmp_path, 400), signal_at=3, sig=signal.SIGTERM)
    assert j.returncode == 0
    assert "[EXIT HANDLER] Job cancelled, terminating." in j.text
    assert not os.path.exists(j.ckpt_path()) and j.sbatch_calls == []


def test_injected_error_saves_without_resubmit_then_resumes(tmp_path):
    j = Job(tmp_path, 3001, _args(tmp_path, 12, "--raise-error", "--error-step", "6"))
    assert j.returncode == 0
    assert "[EXIT HANDLER] Error during training encountered, saving checkpoint." in j.text
    assert "[EXIT HANDLER] Checkpoint saved at step 6" in j.text
    assert j.sbatch_calls == []
    ck = torch.load(j.ckpt_path(), weights_only=False)
    assert ck["training_step"] == 6 and ck["lr_scheduler"]["last_epoch"] == 6  # exactly 6 updates
    assert all(int(s["step"]) == 6 for s in ck["optimizer"]["state"].values())
    # train.sh always passes --raise-error --error-step; the resumed job must get past that step
    j2 = Job(tmp_path, 3002, _args(tmp_path, 12, "--raise_error", "--error_step", "6", "3001"))
    assert "Training completed" in j2.text and set(j2.losses) == set(range(6, 12))


def test_three_job_usr1_chain_is_bitwise_identical(tmp_path):
    full = Job(tmp_path, 4000, _args(tmp_path, 30, "--save-final", "--dataset-type", "packed"))
    prev, step_target, jid = None, 8, 4001
    losses = {}
    for hop in range(3):
        more = ["--dataset-type", "packed"] + ([str(prev)] if prev else [])
        last = hop == 2
        j = Job(tmp_path, jid, _args(tmp_path, 30, *more, *(["--save-final"] if last else [])),
                signal_at=None if last else step_target)
        assert j.returncode == 0
        losses.update(j.losses)
        prev, jid, step_target = jid, jid + 1, step_target + 8
    assert losses == full.losses
    a = torch.load(full.ckpt_path(), weights_only=False)
    b = torch.load(os.path.join(str(tmp_path), "checkpoints", f"checkpoint_{prev}.ckpt"), weights_only=False)
    assert _digest(a) == _digest(b)
    assert a["data_state"] == b["data_state"]
