"""Fake-Slurm harness: runs train.py as a subprocess with SLURM_JOB_ID / WORKDIR set and
a fake `sbatch` first on PATH that records its argv; sends a signal when a given step shows up in the log."""
import os
import re
import signal
import subprocess
import sys
import threading

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STEP_RE = re.compile(r"Training step: (\d+) \| Loss")
LOSS_RE = re.compile(r"step (\d+): loss ([0-9.]+)")


def make_fake_sbatch(bin_dir: str, record_file: str) -> None:
    os.makedirs(bin_dir, exist_ok=True)
    path = os.path.join(bin_dir, "sbatch")
    with open(path, "w") as f:
        f.write(f"#!/bin/bash\necho \"$@\" >> {record_file}\necho \"Submitted batch job fake\"\n")
    os.chmod(path, 0o755)


class Job:
    def __init__(self, workdir, jobid, args, signal_at=None, sig=signal.SIGUSR1, nproc=1, extra_env=None,
                 timeout=240):
        self.workdir, self.jobid = str(workdir), str(jobid)
        self.bin_dir = os.path.join(self.workdir, "bin")
        self.record = os.path.join(self.workdir, "sbatch_calls.txt")
        make_fake_sbatch(self.bin_dir, self.record)
        env = dict(os.environ)
        env.update({"SLURM_JOB_ID": self.jobid, "WORKDIR": self.workdir,
                    "PATH": self.bin_dir + os.pathsep + env.get("PATH", ""),
                    "PYTHONPATH": ROOT + os.pathsep + env.get("PYTHONPATH", ""), "PYTHONUNBUFFERED": "1"})
        env.pop("FTL_JOB_ID", None)
        if extra_env:
            env.update(extra_env)
        if nproc > 1:
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
                   "--master-addr", "127.0.0.1", "--master-port", str(29600 + os.getpid() % 300),
                   os.path.join(ROOT, "train.py")] + list(args)
        else:
            cmd = [sys.executable, os.path.join(ROOT, "train.py")] + list(args)
        self.lines = []
        self.signalled = False
        self.proc = subprocess.Popen(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                                     cwd=self.workdir, start_new_session=True)

        def pump():
            for line in self.proc.stdout:
                self.lines.append(line.rstrip("\n"))
                m = STEP_RE.search(line)
                if m and signal_at is not None and not self.signalled and int(m.group(1)) >= signal_at:
                    self.signalled = True
                    self._signal_job(sig)

        t = threading.Thread(target=pump, daemon=True)
        t.start()
        try:
            self.returncode = self.proc.wait(timeout=timeout)
        except subprocess.TimeoutExpired:
            os.killpg(self.proc.pid, signal.SIGKILL)
            self.returncode = -9
        t.join(timeout=10)
        self.text = "\n".join(self.lines)

    def _signal_job(self, sig):
        """Like srun's signal forwarding: every task of the job step gets the signal.  Under torchrun the workers
        are signalled directly (the launcher itself neither handles nor forwards SIGUSR1)."""
        import psutil
        try:
            kids = psutil.Process(self.proc.pid).children(recursive=True)
        except psutil.NoSuchProcess:
            kids = []
        workers = [k for k in kids if any("train.py" in part for part in (k.cmdline() or []))]
        targets = [k.pid for k in workers] if workers else [self.proc.pid]
        for pid in targets:
            try:
                os.kill(pid, sig)
            except ProcessLookupError:
                pass

    @property
    def losses(self):
        out = {}
        for line in self.lines:
            m = LOSS_RE.search(line)
            if m:
                out[int(m.group(1))] = m.group(2)
        return out

    @property
    def sbatch_calls(self):
        if not os.path.exists(self.record):
            return []
        return [l.strip() for l in open(self.record) if l.strip()]

    def ckpt_path(self, jobid=None):
        return os.path.join(self.workdir, "checkpoints", f"checkpoint_{jobid or self.jobid}.ckpt")
