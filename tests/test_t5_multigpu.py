"""T5 on real GPUs (skipped on a 1-GPU box): every W > 1 kernel template of csrc/comm.cu and the gather-fused GEMMs.

  * data-parallel and fully-sharded trajectories at W = 2, 4, 8 ranks x batch 1 against ONE rank x batch W
    (tools/dp_check.py, tools/fsdp_check.py: losses, parameters, ranks identical, sharded checkpoint round trip);
  * a rank-local failure: the failing rank raises the abort word, the peers' kernels fall through their device
    barriers, ALL ranks checkpoint the same step, and the resumed job completes;
  * the barrier's wall-clock budget: a rank that never arrives makes the others return status 1 and leave parameters
    and moments untouched;
  * a stress run of the barrier and of the gather-fused GEMM with randomised per-rank delays.

The reference has no multi-process path at all; these are the guard rails of the new data plane."""
import json
import os
import subprocess
import sys

import pytest
import torch

from ft_harness import Job

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NGPU = torch.cuda.device_count() if torch.cuda.is_available() else 0
# world sizes of the trajectory tests: 2 and the largest power of two the box has (FTL_TEST_WORLDS=2,4,8 runs them all;
# every size costs a torchrun start-up per test, and W = 4 adds no code path that W = 2 and W = 8 do not take)
_all = [w for w in (2, 4, 8) if w <= NGPU]
WORLDS = ([int(x) for x in os.environ["FTL_TEST_WORLDS"].split(",") if int(x) <= NGPU] if os.environ.get("FTL_TEST_WORLDS")
          else sorted({w for w in _all if w == 2 or w == max(_all)}))
needs2 = pytest.mark.skipif(NGPU < 2, reason="needs at least 2 GPUs")


def _torchrun(world, script, *args, env=None, timeout=900):
    e = dict(os.environ)
    e["PYTHONPATH"] = ROOT + os.pathsep + e.get("PYTHONPATH", "")
    e.update(env or {})
    port = 29700 + (os.getpid() + world * 7) % 200
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, script)] + list(args)
    r = subprocess.run(cmd, env=e, capture_output=True, text=True, timeout=timeout, cwd=ROOT)
    out = []
    for line in r.stdout.splitlines():
        line = line.strip()
        if line.startswith("{"):
            try:
                out.append(json.loads(line))
            except ValueError:
                pass
    return r, out


@needs2
@pytest.mark.parametrize("world", WORLDS or [2])
def test_dp_tracks_one_gpu_batch_w(world):
    r, out = _torchrun(world, "tools/dp_check.py")
    assert r.returncode == 0 and out, r.stdout[-2000:] + r.stderr[-2000:]
    assert out[0]["ok"] and out[0]["ranks_identical"] and out[0]["world"] == world, out[0]
    assert out[1]["ckpt_tensors_with_moments"] == out[1]["of"]


@needs2
@pytest.mark.parametrize("world", WORLDS or [2])
def test_fsdp_tracks_one_gpu_batch_w(world):
    r, out = _torchrun(world, "tools/fsdp_check.py")
    assert r.returncode == 0 and out, r.stdout[-2000:] + r.stderr[-2000:]
    assert out[0]["ok"] and out[0]["fused_gather_gemms"], out[0]
    assert out[1]["resume_from_shards_equal"] and out[1]["resume_from_consolidated_equal"], out[1]


@needs2
def test_fsdp_reshard_mode_tracks_one_gpu():
    """Two staging slots, backward re-gathers (FTL_FSDP_RESHARD=1): same trajectory."""
    r, out = _torchrun(2, "tools/fsdp_check.py", env={"FTL_FSDP_RESHARD": "1"})
    assert r.returncode == 0 and out and out[0]["ok"], r.stdout[-2000:] + r.stderr[-2000:]


TINY = ["--tokenizer-name-or-path", "offline", "--vocab-size", "8192", "--dim", "1024", "--n-layers", "3", "--n-heads", "8",
        "--n-kv-heads", "4", "--multiple-of", "256", "--sequence-length", "256", "--batch-size", "1", "--synthetic-docs", "64",
        "--learning-rate", "1e-3", "--lr-warmup-steps", "5", "--logging-frequency", "1", "--device", "cuda"]


@needs2
@pytest.mark.parametrize("mode", ["dp", "fsdp"])
def test_rank_local_failure_all_ranks_save_same_step(tmp_path, mode):
    args = TINY + ["--dataset", os.path.join(str(tmp_path), "d.parquet"), "--training-steps", "14", "--parallel", mode]
    j = Job(tmp_path, 8100, args, nproc=2, extra_env={"FTL_INJECT_RANK_ERROR": "1:6"}, timeout=600)
    assert j.returncode == 0, j.text[-3000:]
    assert "injected rank-local failure on rank 1 at step 6" in j.text
    assert "[EXIT HANDLER] Error during training encountered, saving checkpoint." in j.text
    assert "[EXIT HANDLER] Checkpoint saved at step 6" in j.text, j.text[-3000:]
    assert j.sbatch_calls == []
    k = Job(tmp_path, 8101, args + ["--checkpoint-id", "8100"], nproc=2, timeout=600)
    assert k.returncode == 0 and "Resuming training from training_step 6" in k.text and "Training completed" in k.text, k.text[-3000:]


@needs2
def test_comm_stress_barrier_timeout_and_delays():
    world = min(NGPU, 8)
    r, out = _torchrun(world, "tools/comm_stress.py", "--epochs", "10000", "--gemms", "200")
    assert r.returncode == 0 and out, r.stdout[-2000:] + r.stderr[-2000:]
    res = out[-1]
    assert res["barrier_epochs_ok"] and res["gather_gemm_ok"] and res["timeout_status"] == 1 and res["untouched_after_timeout"], res
    assert res["abort_status"] == 2, res
    assert res["sparse_embedding_reduce_ok"], res


def _multicast_ok() -> bool:
    from fault_tolerant_llm_training_b200.ops import native
    try:
        torch.cuda.set_device(0)
        return native.vmm_supported()[1]
    except Exception:  # noqa: BLE001
        return False


@needs2
def test_vmm_heap_dp_tracks_one_gpu():
    """Symmetric heap built with cuMemCreate + fd passing + cuMemMap instead of CUDA IPC; same peer-load kernels."""
    r, out = _torchrun(2, "tools/dp_check.py", env={"FTL_SYMM": "vmm"})
    assert r.returncode == 0 and out and out[0]["ok"] and out[0]["ranks_identical"], r.stdout[-2000:] + r.stderr[-2000:]


@needs2
@pytest.mark.parametrize("world", WORLDS or [2])
def test_nvls_collectives_track_one_gpu(world):
    """FTL_COLLECTIVE=nvls: multimem.ld_reduce reduce-scatter and multimem.st all-gather through a multicast mapping."""
    if not _multicast_ok():
        pytest.skip("multicast objects (NVLS) are not supported on this GPU / driver")
    r, out = _torchrun(world, "tools/dp_check.py", env={"FTL_COLLECTIVE": "nvls"})
    assert r.returncode == 0 and out and out[0]["ok"] and out[0]["ranks_identical"], r.stdout[-2000:] + r.stderr[-2000:]
    r, out = _torchrun(world, "tools/fsdp_check.py", env={"FTL_COLLECTIVE": "nvls"})
    assert r.returncode == 0 and out and out[0]["ok"], r.stdout[-2000:] + r.stderr[-2000:]


@needs2
@pytest.mark.parametrize("mode", ["dp", "fsdp"])
def test_two_rank_usr1_chain_is_bitwise(tmp_path, mode):
    """Kill/resume on two ranks: SIGUSR1 to both -> same step boundary on both -> sharded checkpoint -> resubmit ->
    resume; losses and the final model / optimizer / scheduler / data state are bitwise those of an uninterrupted
    two-rank run (BASELINE.json config 5 in miniature, on the peer-load data plane)."""
    import signal

    from test_t2_fault_tolerance import _digest

    mk = lambda steps, *more: TINY + ["--dataset", os.path.join(str(tmp_path), "d.parquet"), "--training-steps", str(steps),
                                      "--parallel", mode] + list(more)
    # enough steps that the job is still running when the signal arrives (a step of this model takes a few ms)
    full = Job(tmp_path, 9000, mk(600, "--save-final"), nproc=2, timeout=600)
    assert full.returncode == 0 and "Training completed" in full.text, full.text[-3000:]
    j1 = Job(tmp_path, 9001, mk(600), nproc=2, signal_at=12, sig=signal.SIGUSR1, timeout=600)
    assert j1.returncode == 0 and "[EXIT HANDLER] Job timed out, saving checkpoint." in j1.text, j1.text[-3000:]
    assert j1.sbatch_calls == [f"{tmp_path}/train.sh 9001"]
    j2 = Job(tmp_path, 9002, mk(600, "--save-final", "--checkpoint-id", "9001"), nproc=2, timeout=600)
    assert j2.returncode == 0 and "Training completed" in j2.text, j2.text[-3000:]
    merged = dict(j1.losses)
    merged.update(j2.losses)
    assert merged == full.losses
    a = torch.load(full.ckpt_path(), weights_only=False)
    b = torch.load(j2.ckpt_path(), weights_only=False)
    assert _digest(a) == _digest(b)


@needs2
def test_fsdp_full_gradient_buffer_tracks_one_gpu():
    """FTL_FSDP_SHARD_GRADS=0: the round-1 layout with a model-sized gradient buffer per rank."""
    r, out = _torchrun(2, "tools/fsdp_check.py", env={"FTL_FSDP_SHARD_GRADS": "0"})
    assert r.returncode == 0 and out and out[0]["ok"], r.stdout[-2000:] + r.stderr[-2000:]


@needs2
@pytest.mark.parametrize("reshard", ["0", "1"])
def test_fsdp_sharded_gradients_track_one_gpu(reshard):
    """Sharded gradients (the default, set explicitly here): no model-sized gradient buffer (two layer-sized slots reused by the backward, compact
    reduced shards), with resident and with re-gathered parameters: same trajectory, same checkpoint round trip."""
    r, out = _torchrun(2, "tools/fsdp_check.py", env={"FTL_FSDP_SHARD_GRADS": "1", "FTL_FSDP_RESHARD": reshard})
    assert r.returncode == 0 and out and out[0]["ok"], r.stdout[-2000:] + r.stderr[-2000:]
    assert out[1]["resume_from_shards_equal"] and out[1]["resume_from_consolidated_equal"], out[1]
