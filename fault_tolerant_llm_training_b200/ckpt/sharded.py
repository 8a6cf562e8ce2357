"""Sharded, parallel checkpoints for the flat-state engines, plus the consolidated reference-layout file.

Save path (SIGUSR1 / exception; BASELINE.json "fault-tolerance subsystem rebuilt for sharding"):

  1. every rank streams ITS slice of the flat parameter / exp_avg / exp_avg_sq buffers (48 GB / W for the
     8 B model) from the device through pinned staging buffers into `checkpoint_<JOBID>.shards/rank<r>.bin`
     with the native snapshot engine (csrc/ckpt_io.cu), and fsyncs;
  2. after a barrier rank 0 writes `manifest.json` atomically (temp + fsync + rename + directory fsync).
     From this instant the checkpoint is DURABLE and resumable — the job can be resubmitted;
  3. rank 0 then consolidates the shards into `checkpoint_<JOBID>.ckpt`, the reference's `torch.save`
     layout (utils.py:75-80), which the reference's own resume code loads; the shard directory is removed
     afterwards unless `keep_shards`.

Resume prefers the consolidated file and falls back to the shard directory (a job killed while
consolidating loses nothing).  Shards can be read back into a different world size: the manifest maps
flat-buffer ranges to (file, offset).

Replacing an older checkpoint of the same job id (periodic saves, a second signal) is crash-safe: the new shards
are written into `checkpoint_<JOBID>.shards.new`; only when its manifest is durable is a stale consolidated file
of that id unlinked and the directory swapped in (`.shards` -> `.shards.old`, `.shards.new` -> `.shards`).  The
loader looks at all three names and takes the complete one with the largest training_step, so a kill at any point
of that sequence leaves a resumable checkpoint.  Shards that a resubmitted job may be reading are never deleted by
the job that wrote them: the successor removes them once its own first checkpoint is durable.
"""
from __future__ import annotations

import json
import os
import shutil
import time
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from ..ops import native as N
from .layout import FORMAT_VERSION, atomic_torch_save, checkpoint_file

MANIFEST = "manifest.json"
ALIGN = 4096


def shard_dir(checkpoint_path: str, jobid: str) -> str:
    return os.path.join(checkpoint_path, f"checkpoint_{jobid}.shards")


def best_shard_dir(checkpoint_path: str, jobid: str) -> Optional[str]:
    """The complete shard directory (manifest present and readable) of this id with the largest training_step among
    `.shards`, `.shards.new` (a save that was killed before the swap) and `.shards.old` (killed inside the swap)."""
    base, best, best_step = shard_dir(checkpoint_path, jobid), None, -1
    for d in (base, base + ".new", base + ".old"):
        try:
            step = int(read_manifest(d)["training_step"])
        except (OSError, ValueError, KeyError):
            continue
        if step > best_step:
            best, best_step = d, step
    return best


def _fsync_dir(d: str) -> None:
    try:
        fd = os.open(d, os.O_RDONLY)
        try:
            os.fsync(fd)
        finally:
            os.close(fd)
    except OSError:
        pass


def _atomic_json(obj, path: str) -> None:
    tmp = f"{path}.tmp.{os.getpid()}"
    with open(tmp, "w") as f:
        json.dump(obj, f)
        f.flush()
        os.fsync(f.fileno())
    os.replace(tmp, path)
    _fsync_dir(os.path.dirname(path))


class FlatCheckpointer:
    """Bound to a NativeTrainer (single GPU or data parallel)."""

    def __init__(self, trainer, keep_shards: bool = False, consolidate: bool = True, io_threads: int = 4):
        self.t = trainer
        self.keep_shards, self.do_consolidate, self.io_threads = keep_shards, consolidate, io_threads
        self.timings: Dict[str, float] = {}
        self._pending: Optional[Tuple[str, str, str]] = None
        self.predecessor: Optional[Tuple[str, str]] = None   # (checkpoint_path, id) this job resumed from
        N.ckpt_init(256 << 20)  # staging is allocated up front: the exit path must not depend on free memory

    # -- save ---------------------------------------------------------------------------------------------------
    def save(self, checkpoint_path: str, jobid: str, model, optimizer, lr_scheduler, training_step: int, *,
             data_state=None, rank: int = 0, world_size: int = 1, extra=None, logger=None) -> str:
        t = self.t
        env = t.env
        t0 = time.perf_counter()
        torch.cuda.synchronize()   # the step that was running has retired; buffers are at a step boundary
        final = shard_dir(checkpoint_path, jobid)
        d = final + ".new"          # written here; swapped in only after its manifest is durable
        if env.is_main:
            shutil.rmtree(d, ignore_errors=True)
            os.makedirs(d, exist_ok=True)
        env.barrier()
        segs, my_entries, off = [], [], 0
        for name, lo, hi, piece in t.checkpoint_segments():
            nbytes = (hi - lo) * 2
            assert piece.numel() == hi - lo and piece.is_contiguous()
            segs.append((piece.data_ptr(), nbytes, off))
            my_entries.append({"rank": env.rank, "buffer": name, "lo": lo, "hi": hi, "file": f"rank{env.rank}.bin",
                               "offset": off})
            off = (off + nbytes + ALIGN - 1) // ALIGN * ALIGN
        io = N.ckpt_write(os.path.join(d, f"rank{env.rank}.bin"), segs, nthreads=self.io_threads, fsync=True)
        entries = env.gather_objects(my_entries)   # also the barrier: every shard is on disk past this point
        t_shards = time.perf_counter() - t0
        if env.is_main:
            groups = []
            for g in optimizer.param_groups:
                groups.append({k: (list(v) if isinstance(v, tuple) else v) for k, v in g.items() if k != "params"})
            manifest = {
                "format": "ftl-b200-shards", "format_version": FORMAT_VERSION, "world_size": env.world_size,
                "flat_numel": t.flat.numel, "dtype": "bfloat16",
                "layout": [[nm, list(shape), o, n] for nm, shape, o, n in t.flat.layout],
                "segments": [e for per_rank in entries for e in per_rank],
                "training_step": int(training_step),
                "optimizer": {"param_groups": groups, "steps_done": int(optimizer.steps_done)},
                "lr_scheduler": lr_scheduler.state_dict(), "data_state": data_state,
                "meta": dict(extra or {}, saved_unix_time=time.time()),
            }
            _atomic_json(manifest, os.path.join(d, MANIFEST))
            _fsync_dir(d)
            # durable from here on (the loader accepts `.shards.new`).  Now retire what it replaces: first a stale
            # consolidated file of this id (the loader prefers a .ckpt), then swap the directory in.
            stale = checkpoint_file(checkpoint_path, jobid)
            if os.path.exists(stale):
                os.unlink(stale)
            old = final + ".old"
            shutil.rmtree(old, ignore_errors=True)
            if os.path.isdir(final):
                os.rename(final, old)
            os.rename(d, final)
            _fsync_dir(checkpoint_path)
            shutil.rmtree(old, ignore_errors=True)
            self._drop_predecessor(logger)
        env.barrier()
        d = final
        t_durable = time.perf_counter() - t0
        self.timings = {"shards_s": t_shards, "durable_s": t_durable, "d2h_s": io["d2h_s"], "write_s": io["write_s"],
                        "fsync_s": io["fsync_s"], "bytes_this_rank": sum(s[1] for s in segs)}
        if logger is not None and env.is_main:
            gb = sum(s[1] for s in segs) / 1e9
            logger.info(f"[EXIT HANDLER] shards durable in {t_durable:.2f} s ({gb:.2f} GB per rank x {env.world_size}; "
                        f"D2H {io['d2h_s']:.2f} s, write {io['write_s']:.2f} s, fsync {io['fsync_s']:.2f} s)")
        self._pending = (d, checkpoint_file(checkpoint_path, jobid), jobid)
        return d

    def drop_predecessor(self, logger=None) -> None:
        """Public form for the training driver: called on rank 0 when a resumed job ends without writing a checkpoint of
        its own, so that the shards it resumed from do not stay behind for ever."""
        if self.t.env.is_main:
            self._drop_predecessor(logger)

    def _drop_predecessor(self, logger=None) -> None:
        """This job's own checkpoint is durable: the shards of the checkpoint it resumed from are no longer needed
        by anyone, provided the job that wrote them has finished consolidating (its .ckpt exists)."""
        if self.predecessor is None:
            return
        path, rid = self.predecessor
        if os.path.exists(checkpoint_file(path, rid)):
            base = shard_dir(path, rid)
            for dd in (base, base + ".new", base + ".old"):
                shutil.rmtree(dd, ignore_errors=True)
            self.predecessor = None

    # -- consolidate (rank 0, after the resubmit) ---------------------------------------------------------------
    def finalize(self, logger=None, resubmitted: bool = False) -> Optional[str]:
        """resubmitted: a successor job was started and may be reading the shards right now — they are left in place
        (the successor removes them after its own first durable checkpoint)."""
        if self._pending is None:
            return None
        d, out, _ = self._pending
        self._pending = None
        path = None
        if self.t.env.is_main and self.do_consolidate:
            t0 = time.perf_counter()
            path = consolidate(d, out)
            if logger is not None:
                logger.info(f"[EXIT HANDLER] consolidated reference-layout file written in "
                            f"{time.perf_counter() - t0:.2f} s -> {out}")
            if not self.keep_shards and not resubmitted:
                shutil.rmtree(d, ignore_errors=True)
        self.t.env.barrier()
        return path


# ------------------------------------------------------------------------------------------------------------
def read_manifest(d: str) -> dict:
    with open(os.path.join(d, MANIFEST)) as f:
        return json.load(f)


def _intersections(manifest: dict, buffer: str, lo: int, hi: int, allow_gaps: bool = False):
    """(file, file_byte_offset, dst_element_offset, n_elements) pieces covering [lo, hi) of `buffer`."""
    out, covered = [], 0
    for s in manifest["segments"]:
        if s["buffer"] != buffer:
            continue
        a, b = max(lo, s["lo"]), min(hi, s["hi"])
        if a < b:
            out.append((s["file"], s["offset"] + (a - s["lo"]) * 2, a - lo, b - a))
            covered += b - a
    if covered != hi - lo and not allow_gaps:
        raise ValueError(f"shards do not cover {buffer}[{lo}:{hi}] (covered {covered} elements)")
    return out


def read_range_to_device(d: str, manifest: dict, buffer: str, lo: int, hi: int, dst: torch.Tensor, io_threads: int = 4):
    """dst: bf16 device tensor of hi-lo elements."""
    by_file: Dict[str, list] = {}
    for fname, foff, doff, n in _intersections(manifest, buffer, lo, hi):
        by_file.setdefault(fname, []).append((dst.data_ptr() + doff * 2, n * 2, foff))
    for fname, segs in by_file.items():
        N.ckpt_read(os.path.join(d, fname), segs, nthreads=io_threads)


def read_flat_host(d: str, manifest: dict, buffer: str) -> torch.Tensor:
    """Whole flat buffer as a CPU bf16 tensor; zero-copy (mmap) when one file holds all of it."""
    n = manifest["flat_numel"]
    pieces = _intersections(manifest, buffer, 0, n, allow_gaps=True)   # alignment padding between tensors is not stored
    if len(pieces) == 1 and pieces[0][3] == n:
        fname, foff, _, cnt = pieces[0]
        mm = np.memmap(os.path.join(d, fname), dtype=np.uint16, mode="r", offset=foff, shape=(cnt,))
        return torch.from_numpy(mm).view(torch.bfloat16)
    out = torch.zeros(n, dtype=torch.bfloat16)
    raw = out.view(torch.int16).numpy().view(np.uint16)
    for fname, foff, doff, cnt in pieces:
        raw[doff:doff + cnt] = np.memmap(os.path.join(d, fname), dtype=np.uint16, mode="r", offset=foff, shape=(cnt,))
    return out


def assemble_payload(d: str, manifest: Optional[dict] = None) -> dict:
    """The reference's four-key payload (+ data_state, meta) built from a shard directory."""
    import warnings

    m = manifest or read_manifest(d)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")  # read-only memmap -> torch tensor
        flat = {b: read_flat_host(d, m, b) for b in ("p", "m", "v")}
    model_sd, opt_state = {}, {}
    steps = float(m["optimizer"]["steps_done"])
    for i, (name, shape, off, n) in enumerate(m["layout"]):
        model_sd[name] = flat["p"][off:off + n].view(shape)
        opt_state[i] = {"step": torch.tensor(steps, dtype=torch.float32),
                        "exp_avg": flat["m"][off:off + n].view(shape), "exp_avg_sq": flat["v"][off:off + n].view(shape)}
    groups = []
    for g in m["optimizer"]["param_groups"]:
        g = dict(g)
        if "betas" in g:
            g["betas"] = tuple(g["betas"])
        g["params"] = list(range(len(m["layout"])))
        groups.append(g)
    payload = {"model": model_sd, "optimizer": {"state": opt_state, "param_groups": groups},
               "lr_scheduler": m["lr_scheduler"], "training_step": int(m["training_step"])}
    if m.get("data_state") is not None:
        payload["data_state"] = m["data_state"]
    payload["meta"] = dict(m.get("meta") or {}, format_version=FORMAT_VERSION, consolidated_from_shards=m["world_size"])
    return payload


def consolidate(d: str, out_file: str) -> str:
    atomic_torch_save(assemble_payload(d), out_file)
    return out_file


def load_sharded(d: str) -> dict:
    """Light-weight checkpoint dict for engines that read tensors straight from the shard files."""
    m = read_manifest(d)
    return {"_sharded_dir": d, "_manifest": m, "training_step": int(m["training_step"]),
            "lr_scheduler": m["lr_scheduler"], "data_state": m.get("data_state"), "meta": m.get("meta")}
